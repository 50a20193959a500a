// VecTask.step prologue / epilogue for B200 (sm_100a) -- SURVEY 8(f) row N4.
//
// Replaces the non-physical domain-randomisation lambdas and the clamp copies either side of the task hooks in
// /root/reference/isaacgymenvs/tasks/base/vec_task.py:
//   step :323-327   actions = noise_lambda(actions); action_tensor = clamp(actions, +-clip_actions)
//   step :347-353   obs_buf = noise_lambda(obs_buf); obs_dict["obs"] = clamp(obs_buf, +-clip_obs)
//   apply_randomizations :576-646   the gaussian / uniform, additive / scaling lambdas with a correlated component
// In the reference one call of a lambda is ~8 elementwise kernels and as many passes over the tensor (randn_like, mul,
// add, corr mul, corr add, add, op) plus the clamp copy; here it is ONE pass: 128-bit loads of the tensor (+ the
// persistent correlated draw), the noise in registers (supplied unit draws or in-kernel Philox), and up to two 128-bit
// stores: the noisy tensor in place (what `self.obs_buf = ...` rebinds to) and its clamped copy wherever the consumer
// wants it -- e.g. straight into slot t of an rl_games [horizon, N, obs] rollout buffer.
// Algorithmic bytes per element: 4 read (+4 correlated draw, +4 supplied unit draw) + 4 per destination written.
#include "tk_common.cuh"

namespace tk {
namespace epilogue {

struct Lambda {
    int dist;              // 0 none, 1 gaussian, 2 uniform
    int scaling;           // 0: tensor + noise, 1: tensor * noise
    float a, b;            // gaussian: var, mu | uniform: hi - lo, lo
    float a_corr, b_corr;  // gaussian: var_corr, mu_corr | uniform: hi_corr - lo_corr, lo_corr
};

enum { kStreamEpilogue = 64 };

__device__ __forceinline__ float apply_one(float x, const Lambda& l, float corr, float z, float clip, float* noisy) {
    if (l.dist != 0) {
        const float c = corr * l.a_corr + l.b_corr;        // :617 / :639
        const float nz = (c + z * l.a) + l.b;              // :619 / :640  corr + rand * a + b
        x = l.scaling ? x * nz : x + nz;                   // op, :592
    }
    *noisy = x;
    return fminf(fmaxf(x, -clip), clip);                   // :327 / :353
}

// one thread = four consecutive elements.  unit: supplied draws (N(0,1) for gaussian, U(0,1) for uniform) or NULL = Philox
// keyed by (seed, step, element index / 4).  noisy_out / clamped_out may be NULL or alias `in` (obs_buf in place): `in` is
// therefore NOT __restrict__ / read-only-path qualified; each thread reads its four elements before it writes them.
__global__ void __launch_bounds__(256) noise_clamp_kernel(const float* in, long long count, Lambda l,
                                                          const float* __restrict__ corr, const float* __restrict__ unit,
                                                          unsigned long long seed, unsigned long long step,
                                                          const unsigned long long* __restrict__ epoch, long long elem_off,
                                                          float clip, float* noisy_out, float* clamped_out, int vec) {
    const long long i4 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
    if (i4 >= count) return;
    const bool full = vec && i4 + 3 < count;
    float x[4] = {0.f, 0.f, 0.f, 0.f}, c[4] = {0.f, 0.f, 0.f, 0.f}, z[4] = {0.f, 0.f, 0.f, 0.f};
    if (full) {
        const float4 v = *reinterpret_cast<const float4*>(in + i4);
        x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
        if (l.dist != 0 && corr) { const float4 w = *reinterpret_cast<const float4*>(corr + i4); c[0] = w.x; c[1] = w.y; c[2] = w.z; c[3] = w.w; }
        if (l.dist != 0 && unit) { const float4 w = *reinterpret_cast<const float4*>(unit + i4); z[0] = w.x; z[1] = w.y; z[2] = w.z; z[3] = w.w; }
    } else {
        for (int k = 0; k < 4; ++k)
            if (i4 + k < count) {
                x[k] = in[i4 + k];
                if (l.dist != 0 && corr) c[k] = corr[i4 + k];
                if (l.dist != 0 && unit) z[k] = unit[i4 + k];
            }
    }
    if (l.dist != 0 && !unit) {
        const Philox4 r = philox4x32_10(seed, static_cast<unsigned long long>((elem_off + i4) >> 2),
                                        (effective_step(step, epoch) << 8) | kStreamEpilogue);
        if (l.dist == 1) {
            const float2 p0 = normal_pair(r.x, r.y), p1 = normal_pair(r.z, r.w);
            z[0] = p0.x; z[1] = p0.y; z[2] = p1.x; z[3] = p1.y;
        } else {
            z[0] = u01(r.x); z[1] = u01(r.y); z[2] = u01(r.z); z[3] = u01(r.w);
        }
    }
    float ny[4], cl[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) cl[k] = apply_one(x[k], l, c[k], z[k], clip, &ny[k]);
    if (full) {
        if (noisy_out) *reinterpret_cast<float4*>(noisy_out + i4) = make_float4(ny[0], ny[1], ny[2], ny[3]);
        if (clamped_out) *reinterpret_cast<float4*>(clamped_out + i4) = make_float4(cl[0], cl[1], cl[2], cl[3]);
    } else {
        for (int k = 0; k < 4; ++k)
            if (i4 + k < count) {
                if (noisy_out) noisy_out[i4 + k] = ny[k];
                if (clamped_out) clamped_out[i4 + k] = cl[k];
            }
    }
}

// the draws noise_clamp_kernel makes in Philox mode (parity checks)
__global__ void philox_unit_kernel(long long count, int dist, unsigned long long seed, unsigned long long step,
                                   const unsigned long long* __restrict__ epoch, long long elem_off, float* __restrict__ out) {
    const long long i4 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
    if (i4 >= count) return;
    const Philox4 r = philox4x32_10(seed, static_cast<unsigned long long>((elem_off + i4) >> 2),
                                    (effective_step(step, epoch) << 8) | kStreamEpilogue);
    float z[4];
    if (dist == 1) {
        const float2 p0 = normal_pair(r.x, r.y), p1 = normal_pair(r.z, r.w);
        z[0] = p0.x; z[1] = p0.y; z[2] = p1.x; z[3] = p1.y;
    } else {
        z[0] = u01(r.x); z[1] = u01(r.y); z[2] = u01(r.z); z[3] = u01(r.w);
    }
    for (int k = 0; k < 4; ++k)
        if (i4 + k < count) out[i4 + k] = z[k];
}

__device__ unsigned long long g_zero_epoch = 0ull;

}  // namespace epilogue
}  // namespace tk

using namespace tk;
using namespace tk::epilogue;

static int to_lambda(const TkNoiseLambda* nl, Lambda& l) {
    l = Lambda{0, 0, 0.f, 0.f, 0.f, 0.f};
    if (!nl || nl->distribution == 0) return TK_OK;
    TK_REQUIRE(nl->distribution == 1 || nl->distribution == 2, TK_ERR_INVALID, "distribution %d: 0 none, 1 gaussian, 2 uniform",
               nl->distribution);
    TK_REQUIRE(nl->operation == 0 || nl->operation == 1, TK_ERR_INVALID, "operation %d: 0 additive, 1 scaling", nl->operation);
    l.dist = nl->distribution; l.scaling = nl->operation;
    if (nl->distribution == 1) {          // gaussian: range = [mu, var], vec_task.py:591
        l.a = static_cast<float>(nl->range[1]); l.b = static_cast<float>(nl->range[0]);
        l.a_corr = static_cast<float>(nl->range_correlated[1]); l.b_corr = static_cast<float>(nl->range_correlated[0]);
    } else {                              // uniform: range = [lo, hi], the spans are formed in double, :639-640
        l.a = static_cast<float>(nl->range[1] - nl->range[0]); l.b = static_cast<float>(nl->range[0]);
        l.a_corr = static_cast<float>(nl->range_correlated[1] - nl->range_correlated[0]);
        l.b_corr = static_cast<float>(nl->range_correlated[0]);
    }
    return TK_OK;
}

// ---------------------------------------------------------------------------------------------
// Reset id list: env_ids = reset_buf.nonzero().squeeze(-1) -> int32 (humanoid.py:290 + :264, gogoro_new.py:377 + :506),
// ascending like nonzero().  Two small launches: per-chunk counts, then each chunk sums the counts in front of it
// (<= 1024 ints at 1M envs) and writes its ids in order.  4 consecutive envs per thread, 1024 per block.
// ---------------------------------------------------------------------------------------------
constexpr int kIdsThreads = 256;
constexpr int kIdsPerThread = 4;
constexpr int kIdsChunk = kIdsThreads * kIdsPerThread;

__device__ __forceinline__ unsigned ids_flags(const long long* __restrict__ reset, long long n, long long first) {
    unsigned m = 0;
#pragma unroll
    for (int j = 0; j < kIdsPerThread; ++j)
        if (first + j < n && reset[first + j] != 0) m |= 1u << j;
    return m;
}

// block-wide sum of one int per thread -> every thread; red[kIdsThreads / 32 + 1] shared
__device__ __forceinline__ int ids_block_sum(int v, int* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();                      // red may still be read by a previous use
    if (lane == 0) red[warp] = v;
    __syncthreads();
    int t = 0;
#pragma unroll
    for (int w = 0; w < kIdsThreads / 32; ++w) t += red[w];
    return t;
}

__global__ void __launch_bounds__(kIdsThreads) reset_ids_count_kernel(const long long* __restrict__ reset, long long n,
                                                                      int* __restrict__ chunk_counts) {
    __shared__ int red[kIdsThreads / 32];
    const long long first = (static_cast<long long>(blockIdx.x) * kIdsThreads + threadIdx.x) * kIdsPerThread;
    const int total = ids_block_sum(__popc(ids_flags(reset, n, first)), red);
    if (threadIdx.x == 0) chunk_counts[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kIdsThreads) reset_ids_write_kernel(const long long* __restrict__ reset, long long n,
                                                                      const int* __restrict__ chunk_counts, int* __restrict__ ids_out,
                                                                      int* __restrict__ count_out) {
    __shared__ int red[kIdsThreads / 32];
    __shared__ int warp_excl[kIdsThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int before = 0;
    for (int j = threadIdx.x; j < static_cast<int>(blockIdx.x); j += kIdsThreads) before += chunk_counts[j];
    const int base = ids_block_sum(before, red);
    const long long first = (static_cast<long long>(blockIdx.x) * kIdsThreads + threadIdx.x) * kIdsPerThread;
    const unsigned m = ids_flags(reset, n, first);
    const int c = __popc(m);
    int incl = c;                          // inclusive scan inside the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    __syncthreads();
    if (lane == 31) red[warp] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
        int run = 0;
#pragma unroll
        for (int w = 0; w < kIdsThreads / 32; ++w) { warp_excl[w] = run; run += red[w]; }
        if (blockIdx.x == gridDim.x - 1) *count_out = base + run;
    }
    __syncthreads();
    int at = base + warp_excl[warp] + incl - c;
#pragma unroll
    for (int j = 0; j < kIdsPerThread; ++j)
        if (m & (1u << j)) ids_out[at++] = static_cast<int>(first + j);
}

extern "C" {

int tk_step_noise_clamp(const float* in, int64_t count, const TkNoiseLambda* lambda, const float* corr, const float* unit,
                        uint64_t seed, uint64_t step, int64_t elem_offset, float clip, float* noisy_out, float* clamped_out,
                        void* stream) {
    TK_REQUIRE(in != nullptr && count >= 0, TK_ERR_INVALID, "bad input tensor");
    TK_REQUIRE(noisy_out || clamped_out, TK_ERR_INVALID, "no destination");
    if (count == 0) return TK_OK;
    Lambda l;
    if (int rc = to_lambda(lambda, l)) return rc;
    TK_REQUIRE(!(clip < 0.0f), TK_ERR_INVALID, "negative clip");
    const int vec = aligned_to(in, 16) && (!corr || aligned_to(corr, 16)) && (!unit || aligned_to(unit, 16)) &&
                    (!noisy_out || aligned_to(noisy_out, 16)) && (!clamped_out || aligned_to(clamped_out, 16));
    unsigned long long* zero_epoch = nullptr;
    TK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&zero_epoch), g_zero_epoch));
    const long long threads = (count + 3) / 4;
    noise_clamp_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        in, count, l, corr, unit, seed, step, zero_epoch, elem_offset, clip, noisy_out, clamped_out, vec);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_step_noise_units(int64_t count, int32_t distribution, uint64_t seed, uint64_t step, int64_t elem_offset, float* unit_out,
                        void* stream) {
    TK_REQUIRE(unit_out != nullptr && count >= 0 && (distribution == 1 || distribution == 2), TK_ERR_INVALID, "bad arguments");
    if (count == 0) return TK_OK;
    unsigned long long* zero_epoch = nullptr;
    TK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&zero_epoch), g_zero_epoch));
    const long long threads = (count + 3) / 4;
    philox_unit_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        count, distribution, seed, step, zero_epoch, elem_offset, unit_out);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int64_t tk_reset_ids_scratch_bytes(int64_t num_envs) {
    if (num_envs <= 0) return 0;
    return static_cast<int64_t>(sizeof(int)) * ((num_envs + kIdsChunk - 1) / kIdsChunk);
}

int tk_reset_ids(const int64_t* reset_buf, int64_t num_envs, int32_t* ids_out, int32_t* count_out, int32_t* scratch, void* stream) {
    TK_REQUIRE(reset_buf && ids_out && count_out && scratch, TK_ERR_INVALID, "null argument");
    TK_REQUIRE(num_envs > 0 && num_envs <= 2147483647ll, TK_ERR_INVALID, "num_envs %lld does not fit int32 env ids", static_cast<long long>(num_envs));
    TK_REQUIRE(aligned_to(reset_buf, 8) && aligned_to(ids_out, 4) && aligned_to(count_out, 4) && aligned_to(scratch, 4), TK_ERR_ALIGN,
               "reset_buf must be 8-byte, the int32 arrays 4-byte aligned");
    const unsigned chunks = static_cast<unsigned>((num_envs + kIdsChunk - 1) / kIdsChunk);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long* reset = reinterpret_cast<const long long*>(reset_buf);
    reset_ids_count_kernel<<<chunks, kIdsThreads, 0, st>>>(reset, num_envs, scratch);
    reset_ids_write_kernel<<<chunks, kIdsThreads, 0, st>>>(reset, num_envs, scratch, ids_out, count_out);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

}  // extern "C"
