import sys,json
d=json.loads(sys.stdin.readlines()[-1]); r=d["roofline"]; o=r["other_kernels"]
print("k",r["sweeps_per_launch"],"value",round(d["value"],3),"ms/step",round(d["ms_per_step"],2),"launch_ms",round(r["launch_ms"],3),"Gcs/s",round(r["gcell_sweeps_per_s"],1),"pred",round(o["predictor_ms"],2),"corr",round(o["corrector_ms"],2))
