// Developer probe: how long does the FIRST global load of a CTA take after griddepcontrol.wait releases?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe_first_load.bin tools/probe_first_load.cu
// Chain of PDL launches, grid x 128 threads; each kernel: [optional L2 prefetch] -> griddepcontrol.wait -> clock/globaltimer
// -> one coalesced 4-byte load per thread from `arrays` different allocations -> stamp -> small store.  Reports, per
// variant, the median / max cycles from the wait to the landed load and the launch-to-launch time.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

struct P { const float* a[8]; float* out; long long* ts; int arrays; int prefetch; int work; int rare; };

__device__ __noinline__ void rare_block(float& a0, float& a1, float& a2, float& a3, float& a4, float& a5, float& a6, float& a7) {
#pragma unroll
    for (int j = 0; j < 64; ++j) {
        a0 = a0 * 1.0011f + 0.25f; a1 = a1 * 1.0012f + 0.25f; a2 = a2 * 1.0013f + 0.25f; a3 = a3 * 1.0014f + 0.25f;
        a4 = a4 * 1.0015f + 0.25f; a5 = a5 * 1.0016f + 0.25f; a6 = a6 * 1.0017f + 0.25f; a7 = a7 * 1.0018f + 0.25f;
    }
}

__global__ void __launch_bounds__(128) probe(const __grid_constant__ P p) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (p.prefetch && threadIdx.x < p.arrays)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p.a[threadIdx.x] + blockIdx.x * 128));
    if (p.rare == 2 && threadIdx.x < 32) {        // instruction warm-up: the rare block on dummy values, ahead of the wait
        float d0 = i, d1 = 1, d2 = 2, d3 = 3, d4 = 4, d5 = 5, d6 = 6, d7 = 7;
        rare_block(d0, d1, d2, d3, d4, d5, d6, d7);
        if (d0 + d1 + d2 + d3 + d4 + d5 + d6 + d7 == 123.456f) p.out[1] = d0;
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    long long g0, g1;
    const long long c0 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
    float s = 0.0f;
    for (int k = 0; k < p.arrays; ++k) s += p.a[k][i];
    if (s == 123.456f) p.out[0] = s;                       // consumes the loads
    const long long c1 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
    // straight-line block of 2048 FFMAs (8 independent chains, 32 KB of code), executed p.work times: pass 0 runs from
    // whatever the instruction caches hold after the kernel boundary, pass 1 from a warm cache
    float a0 = s, a1 = s + 1, a2 = s + 2, a3 = s + 3, a4 = s + 4, a5 = s + 5, a6 = s + 6, a7 = s + 7;
    long long pass_cyc[2] = {0, 0};
#pragma unroll 1
    for (int k = 0; k < p.work; ++k) {
        const long long t0 = clock64();
#pragma unroll
        for (int j = 0; j < 256; ++j) {
            a0 = a0 * 1.0001f + 0.5f; a1 = a1 * 1.0002f + 0.5f; a2 = a2 * 1.0003f + 0.5f; a3 = a3 * 1.0004f + 0.5f;
            a4 = a4 * 1.0005f + 0.5f; a5 = a5 * 1.0006f + 0.5f; a6 = a6 * 1.0007f + 0.5f; a7 = a7 * 1.0008f + 0.5f;
        }
        if (k < 2) pass_cyc[k] = clock64() - t0;
    }
    // a RARE path: one warp of every 9th CTA runs a 16 KB straight-line block twice.  pass 0 = first execution of that code
    // on this SM in this launch, pass 1 = warm.  (p.rare = 2: every CTA ran the same block once BEFORE the wait.)
    long long rare_cyc[2] = {0, 0};
    if (p.rare && blockIdx.x % 9 == 0 && threadIdx.x < 32) {
#pragma unroll 1
        for (int k = 0; k < 2; ++k) {
            const long long t0 = clock64();
            rare_block(a0, a1, a2, a3, a4, a5, a6, a7);
            rare_cyc[k] = clock64() - t0;
        }
    }
    p.out[i] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (threadIdx.x == 0 && p.ts && p.rare) { p.ts[blockIdx.x * 4 + 3] = (rare_cyc[0] << 32) | rare_cyc[1]; }
    if (threadIdx.x == 0 && p.ts) { p.ts[blockIdx.x * 4 + 0] = c1 - c0; p.ts[blockIdx.x * 4 + 1] = g0; p.ts[blockIdx.x * 4 + 2] = g1; if (!p.rare) p.ts[blockIdx.x * 4 + 3] = (pass_cyc[0] << 32) | pass_cyc[1]; }
}

static void launch(const P& p, int grid, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(128); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, probe, p);
}

int main() {
    const int grid = 512;
    cudaStream_t st; cudaStreamCreate(&st);
    for (int sets : {2, 40}) for (int arrays : {1, 8}) for (int prefetch : {0, 1}) for (int rare : {0, 1, 2}) {
        const int work = 0;
        if (rare && !(arrays == 8 && prefetch == 1)) continue;
        // each "set" = 8 input arrays + 1 output, each in its own 32 MB allocation (different pages; sets x 9 x 32 MB total)
        std::vector<P> ps(sets);
        std::vector<void*> allocs;
        for (int s = 0; s < sets; ++s) {
            P p = {};
            for (int k = 0; k < 8; ++k) { void* d; cudaMalloc(&d, 32 << 20); cudaMemset(d, 0, 32 << 20); allocs.push_back(d); p.a[k] = (const float*)d + (s * 977 % 4096) * 128; }
            void* o; cudaMalloc(&o, 32 << 20); allocs.push_back(o); p.out = (float*)o;
            p.ts = nullptr; p.arrays = arrays; p.prefetch = prefetch; p.work = work; p.rare = rare;
            ps[s] = p;
        }
        long long* ts; cudaMalloc(&ts, grid * 4 * sizeof(long long));
        for (int i = 0; i < 3 * sets; ++i) launch(ps[i % sets], grid, st);
        cudaStreamSynchronize(st);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        const int reps = 400;
        cudaEventRecord(e0, st);
        for (int i = 0; i < reps; ++i) launch(ps[i % sets], grid, st);
        cudaEventRecord(e1, st);
        cudaStreamSynchronize(st);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        // probed launch in the middle of a chain
        for (int i = 0; i < sets; ++i) launch(ps[i % sets], grid, st);
        P pp = ps[1 % sets]; pp.ts = ts;
        launch(ps[0], grid, st);
        launch(pp, grid, st);
        launch(ps[0], grid, st);
        cudaStreamSynchronize(st);
        std::vector<long long> h(grid * 4);
        cudaMemcpy(h.data(), ts, grid * 4 * sizeof(long long), cudaMemcpyDeviceToHost);
        std::vector<long long> cyc, gt;
        long long gmin = h[1];
        for (int b = 0; b < grid; ++b) { cyc.push_back(h[b * 4]); gt.push_back(h[b * 4 + 2] - h[b * 4 + 1]); gmin = std::min(gmin, h[b * 4 + 1]); }
        long long glast = 0;
        for (int b = 0; b < grid; ++b) glast = std::max(glast, h[b * 4 + 2] - gmin);
        std::vector<long long> w0, w1;
        for (int b = 0; b < grid; ++b) { w0.push_back(h[b * 4 + 3] >> 32); w1.push_back(h[b * 4 + 3] & 0xffffffffll); }
        std::sort(w0.begin(), w0.end()); std::sort(w1.begin(), w1.end());
        if (rare) {
            std::vector<long long> r0, r1;
            for (int b = 0; b < grid; b += 9) { r0.push_back(h[b * 4 + 3] >> 32); r1.push_back(h[b * 4 + 3] & 0xffffffffll); }
            std::sort(r0.begin(), r0.end()); std::sort(r1.begin(), r1.end());
            printf("{\"rare\": %d, \"rare_block_cycles\": {\"pass0_min\": %lld, \"pass0_median\": %lld, \"pass0_max\": %lld, \"pass1_median\": %lld, \"pass1_max\": %lld}}\n", rare,
                   r0.front(), r0[r0.size() / 2], r0.back(), r1[r1.size() / 2], r1.back());
        }
        if (work) printf("{\"code_block_2048_ffma_cycles\": {\"pass0_median\": %lld, \"pass0_max\": %lld, \"pass1_median\": %lld, \"pass1_max\": %lld}}\n", w0[grid / 2], w0.back(), w1[grid / 2], w1.back());
        std::sort(cyc.begin(), cyc.end()); std::sort(gt.begin(), gt.end());
        printf("{\"sets\": %d, \"arrays\": %d, \"prefetch\": %d, \"us_per_launch\": %.2f, \"wait_to_load_cycles\": {\"min\": %lld, \"median\": %lld, \"max\": %lld}, "
               "\"wait_to_load_ns_globaltimer_median\": %lld, \"first_wait_to_last_load_ns\": %lld}\n",
               sets, arrays, prefetch, 1e3 * ms / reps, cyc.front(), cyc[grid / 2], cyc.back(), gt[grid / 2], glast);
        fflush(stdout);
        for (void* d : allocs) cudaFree(d);
        cudaFree(ts);
    }
    return 0;
}
