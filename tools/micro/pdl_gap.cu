// pdl_gap.cu — how long is the hole between two dependent kernels of one stream on this box, with and without
// programmatic dependent launch?  Each kernel stamps %globaltimer when its first CTA starts, when its first CTA gets
// past griddepcontrol.wait, and when its last CTA ends.  usage: pdl_gap [ctas] [spin_ns] [alternate 24 KB / 2 KB static shared memory 0|1] [carve-out % for every kernel, -1 = default]
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
struct Stamps { unsigned long long start_inv, go_inv, end; };
struct Big { char pad[256]; };
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
template <bool EARLY, int SMEM>
__global__ void k(Stamps* s, int slot, unsigned spin_ns, Big b) {
    __shared__ int pad[SMEM / 4];
    if (spin_ns == 1u) pad[threadIdx.x] = slot;
    if (spin_ns == 2u) s[1].end = pad[(threadIdx.x + 7) & 255];
    if (threadIdx.x == 0) atomicMax(&s[slot].start_inv, ~gtime());
    if (EARLY) asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (!EARLY) asm volatile("griddepcontrol.launch_dependents;");
    if (threadIdx.x == 0) atomicMax(&s[slot].go_inv, ~gtime());
    const unsigned long long t0 = gtime();
    while (gtime() - t0 < spin_ns) {}
    if (threadIdx.x == 0) atomicMax(&s[slot].end, gtime());
    if (b.pad[0] == 77) s[0].end = 0;
}
int main(int argc, char** argv) {
    const int ctas = argc > 1 ? atoi(argv[1]) : 148 * 6, spin = argc > 2 ? atoi(argv[2]) : 5000, n = 64;
    const int alt = argc > 3 ? atoi(argv[3]) : 0, carve = argc > 4 ? atoi(argv[4]) : -1;
    if (carve >= 0) {
        cudaFuncSetAttribute(k<true, 24576>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        cudaFuncSetAttribute(k<true, 2048>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        cudaFuncSetAttribute(k<false, 24576>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        cudaFuncSetAttribute(k<false, 2048>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
    }
    Stamps* d; cudaMalloc(&d, n * sizeof(Stamps));
    Stamps* h = (Stamps*)malloc(n * sizeof(Stamps));
    cudaStream_t st; cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    Big b = {};
    for (int mode = 0; mode < 3; ++mode) {       // 0 plain, 1 PDL (launch_dependents after wait), 2 PDL (before wait)
        for (int rep = 0; rep < 2; ++rep) {
            cudaMemsetAsync(d, 0, n * sizeof(Stamps), st);
            for (int i = 0; i < n; ++i) {
                cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(256); cfg.stream = st;
                cudaLaunchAttribute a[1]; a[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; a[0].val.programmaticStreamSerializationAllowed = 1;
                cfg.attrs = a; cfg.numAttrs = mode ? 1 : 0;
                // odd kernels use 24 KB of static shared memory per CTA, even ones 2 KB (k_setup / k_tile of the renderer)
                const bool fat = alt && (i & 1);
                if (mode == 2) { if (fat) cudaLaunchKernelEx(&cfg, k<true, 24576>, d, i, (unsigned)spin, b); else cudaLaunchKernelEx(&cfg, k<true, 2048>, d, i, (unsigned)spin, b); }
                else { if (fat) cudaLaunchKernelEx(&cfg, k<false, 24576>, d, i, (unsigned)spin, b); else cudaLaunchKernelEx(&cfg, k<false, 2048>, d, i, (unsigned)spin, b); }
            }
            cudaStreamSynchronize(st);
        }
        cudaMemcpy(h, d, n * sizeof(Stamps), cudaMemcpyDeviceToHost);
        double gap = 0, wait = 0, dur = 0;
        for (int i = 9; i < n; ++i) { gap += (double)(~h[i].start_inv) - (double)h[i - 1].end; wait += (double)(~h[i].go_inv) - (double)h[i - 1].end; dur += (double)h[i].end - (double)(~h[i].go_inv); }
        printf("alt-smem %d carveout %d | mode %d (%s) ctas %d spin %d ns: first CTA starts %.2f us after previous kernel's end, passes the dependency %.2f us after it; body %.2f us; err %s\n",
               alt, carve, mode, mode == 0 ? "plain" : mode == 1 ? "PDL, trigger after wait" : "PDL, trigger before wait", ctas, spin,
               gap / (n - 9) * 1e-3, wait / (n - 9) * 1e-3, dur / (n - 9) * 1e-3, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
