// Micro-benchmark: what one neighbour exchange through L2 costs, with nothing else going on.
// CTAs form a chain; in every round each thread stores 32 bytes of {data, round} packets for the CTA above and
// for the CTA below, then spins until the packets of both neighbours carry the round's number (the protocol of
// csrc/resident_kernels.cuh). Prints cycles per round for several shapes and several kinds of load / store.
// Every spin is bounded, so a flavour that never sees the data reports "stuck" instead of hanging the GPU.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/l2_exchange tools/micro/l2_exchange.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

enum { SYS_SYS = 0, GPU_GPU = 1, WEAKST_GPULD = 2, WEAKST_SYSLD = 3, GPUST_ASYNC = 4, NARROW_GPU = 5 };
static const char* kNames[] = {"st.volatile + ld.volatile (system scope)", "st.relaxed.gpu + ld.relaxed.gpu", "plain st + ld.relaxed.gpu",
                               "plain st + ld.volatile", "st.relaxed.gpu + cp.async.cg into shared memory", "8-byte st/ld.relaxed.gpu (four per 32 bytes)"};

template <int F> __device__ __forceinline__ void put32(unsigned long long* dst, unsigned long long v) {
    if (F == SYS_SYS) {
        asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(v), "l"(v) : "memory");
        asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst + 2), "l"(v), "l"(v) : "memory");
    } else if (F == GPU_GPU || F == GPUST_ASYNC) {
        asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(v), "l"(v) : "memory");
        asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(dst + 2), "l"(v), "l"(v) : "memory");
    } else if (F == NARROW_GPU) {
        for (int k = 0; k < 4; ++k) asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(dst + k), "l"(v) : "memory");
    } else {
        asm volatile("st.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(v), "l"(v) : "memory");
        asm volatile("st.global.v2.u64 [%0], {%1, %2};" ::"l"(dst + 2), "l"(v), "l"(v) : "memory");
    }
}
template <int F> __device__ __forceinline__ bool get32(const unsigned long long* src, unsigned r, unsigned char* slot) {
    unsigned long long a, b, c, d;
    if (F == SYS_SYS || F == WEAKST_SYSLD) {
        asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(src) : "memory");
        asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(c), "=l"(d) : "l"(src + 2) : "memory");
    } else if (F == GPUST_ASYNC) {
        const unsigned s = (unsigned)__cvta_generic_to_shared(slot);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(src) : "memory");
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s + 16), "l"(src + 2) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        const ulonglong2 q0 = *reinterpret_cast<const ulonglong2*>(slot), q1 = *reinterpret_cast<const ulonglong2*>(slot + 16);
        a = q0.x; b = q0.y; c = q1.x; d = q1.y;
    } else if (F == NARROW_GPU) {
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(a) : "l"(src) : "memory");
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(b) : "l"(src + 1) : "memory");
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(c) : "l"(src + 2) : "memory");
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(d) : "l"(src + 3) : "memory");
    } else {
        asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(src) : "memory");
        asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(c), "=l"(d) : "l"(src + 2) : "memory");
    }
    return (a >> 32) == r && (b >> 32) == r && (c >> 32) == r && (d >> 32) == r;
}

template <int F> __global__ void exchange(unsigned long long* band, int rounds, int filler, long long* out) {
    __shared__ __align__(16) unsigned char slots[256 * 64];
    const int tid = threadIdx.x, nt = blockDim.x, cta = blockIdx.x, n = gridDim.x;
    // band layout: [parity][cta][side: 0 = read by the CTA above, 1 = read by the CTA below][thread][4 packets]
    auto slot = [&](int par, int c, int side) { return band + ((((size_t)par * n + c) * 2 + side) * nt + tid) * 4; };
    const bool up = cta > 0, dn = cta < n - 1;
    float acc = (float)tid;
    bool stuck = false;
    long long polls = 0;
    const long long t0 = clock64();
    for (int r = 1; r <= rounds && !stuck; ++r) {
        const int par = r & 1;
        const unsigned long long v = ((unsigned long long)r << 32) | (unsigned)tid;
        if (up) put32<F>(slot(par, cta, 0), v);
        if (dn) put32<F>(slot(par, cta, 1), v);
        for (int k = 0; k < filler; ++k) acc = __fmaf_rn(acc, 1.0000001f, 0.5f);   // stand-in for the rows between
        bool ok;
        int spins = 0;
        do {
            ok = true;
            if (up) ok = get32<F>(slot(par, cta - 1, 1), r, slots + tid * 64) && ok;
            if (dn) ok = get32<F>(slot(par, cta + 1, 0), r, slots + tid * 64 + 32) && ok;
            ++polls;
            if (++spins > (1 << 18)) { stuck = true; break; }
        } while (!ok);
        if (__syncthreads_or(stuck)) stuck = true;
        if (stuck) atomicExch((unsigned long long*)(out + 2), 1ull);
        if (*((volatile long long*)(out + 2))) break;   // (somebody gave up: everybody leaves)
    }
    const long long t1 = clock64();
    if (tid == 0 && cta == n / 2) { out[0] = t1 - t0; out[1] = polls; }
    if (acc == 12345.f) out[3] = 1;
}

template <int F> void run(unsigned long long* band, long long* out, int ctas, int threads, int filler) {
    long long h[4];
    cudaMemset(band, 0, 64 << 20);
    cudaMemset(out, 0, 32);
    const int rounds = 2000;
    void* args[] = {&band, (void*)&rounds, (void*)&filler, &out};
    cudaError_t e = cudaLaunchCooperativeKernel((void*)exchange<F>, dim3(ctas), dim3(threads), args, 0, 0);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("failed: %s\n", cudaGetErrorString(e)); exit(1); }
    cudaMemcpy(h, out, 32, cudaMemcpyDeviceToHost);
    printf("  %-50s %3d CTAs x %3d threads, %3d FMAs between: ", kNames[F], ctas, threads, filler);
    if (h[2]) printf("STUCK (the data never showed up)\n");
    else printf("%6.0f cycles per round, %.2f polls per round\n", (double)h[0] / rounds, (double)h[1] / rounds);
    fflush(stdout);
}

int main() {
    unsigned long long* band;
    long long* out;
    cudaMalloc(&band, 64 << 20);
    cudaMalloc(&out, 32);
    const int shapes[][3] = {{2, 32, 0}, {2, 256, 0}, {125, 32, 0}, {125, 128, 0}, {125, 256, 0}, {125, 256, 400}};
    for (auto& s : shapes) {
        run<SYS_SYS>(band, out, s[0], s[1], s[2]);
        run<GPU_GPU>(band, out, s[0], s[1], s[2]);
        run<WEAKST_GPULD>(band, out, s[0], s[1], s[2]);
        run<WEAKST_SYSLD>(band, out, s[0], s[1], s[2]);
        run<GPUST_ASYNC>(band, out, s[0], s[1], s[2]);
        run<NARROW_GPU>(band, out, s[0], s[1], s[2]);
    }
    return 0;
}
