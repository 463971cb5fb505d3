// native_loop.cu — the timed loop of bench.py without Python or torch in the process: sr_render_device in a loop on
// this program's own stream, L2 flushed between frames, CUDA events per frame.  Used to tell what the host process
// contributes to the holes between a frame's kernels.
//   native_loop <lib.so> <mesh.bin> W H frames [host lag us]     mesh.bin: u64 nverts, u64 nidx, float32[nverts*8], int64[nidx], u64 nmats, float32[nmats*16] (column-major)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <dlfcn.h>
#include <unistd.h>
#include <cuda_runtime.h>
#include "../../include/sr_cuda.h"
int main(int argc, char** argv) {
    if (argc < 6) return 2;
    void* h = dlopen(argv[1], RTLD_NOW);
    if (!h) { fprintf(stderr, "%s\n", dlerror()); return 3; }
#define SYM(n) auto p_##n = reinterpret_cast<decltype(&n)>(dlsym(h, #n))
    SYM(sr_create); SYM(sr_set_mesh); SYM(sr_render_device); SYM(sr_sync); SYM(sr_destroy); SYM(sr_get_stats);
    FILE* f = fopen(argv[2], "rb");
    unsigned long long nv, ni, nm;
    if (!f || fread(&nv, 8, 1, f) != 1 || fread(&ni, 8, 1, f) != 1) return 4;
    std::vector<float> v(nv * 8); std::vector<long long> idx(ni);
    if (fread(v.data(), 4, v.size(), f) != v.size() || fread(idx.data(), 8, ni, f) != ni || fread(&nm, 8, 1, f) != 1) return 4;
    std::vector<float> mats(nm * 16);
    if (fread(mats.data(), 4, mats.size(), f) != mats.size()) return 4;
    const int W = atoi(argv[3]), H = atoi(argv[4]), frames = atoi(argv[5]);
    sr_ctx* c = nullptr;
    if (p_sr_create(&c, 0)) return 5;
    if (p_sr_set_mesh(c, (const sr_vertex*)v.data(), nv, (const int64_t*)idx.data(), ni)) return 6;
    cudaStream_t st; cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    void *color, *depth, *flush;
    cudaMalloc(&color, (size_t)W * H * 4); cudaMalloc(&depth, (size_t)W * H * 4); cudaMalloc(&flush, 256u << 20);
    for (int k = 0; k < 6; ++k) { p_sr_render_device(c, &mats[16 * (k % nm)], 1, SR_TRIANGLE, W, H, color, depth, 0, st); p_sr_sync(c); }
    std::vector<cudaEvent_t> e0(frames), e1(frames);
    for (int k = 0; k < frames; ++k) { cudaEventCreate(&e0[k]); cudaEventCreate(&e1[k]); }
    const int lag_us = argc > 6 ? atoi(argv[6]) : 0;   // host-bound emulation: sleep before every frame's submission
    for (int k = 0; k < frames; ++k) {
        if (lag_us) { cudaStreamSynchronize(st); usleep(lag_us); }
        cudaMemsetAsync(flush, k & 0xff, 256u << 20, st);
        cudaEventRecord(e0[k], st);
        p_sr_render_device(c, &mats[16 * ((6 + k) % nm)], 1, SR_TRIANGLE, W, H, color, depth, 0, st);
        cudaEventRecord(e1[k], st);
    }
    p_sr_sync(c);
    cudaStreamSynchronize(st);
    double tot = 0;
    for (int k = 0; k < frames; ++k) { float ms; cudaEventElapsedTime(&ms, e0[k], e1[k]); tot += ms; }
    sr_stats s; p_sr_get_stats(c, &s);
    printf("native loop: %d frames, %.1f us per frame, %.1f frames/s, launches %llu\n", frames, 1000.0 * tot / frames, frames / (tot / 1000.0), (unsigned long long)s.gpu_launches);
    p_sr_destroy(c);
    return 0;
}
