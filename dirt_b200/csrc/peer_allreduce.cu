// peer_allreduce.cu -- the sum over the ranks of the packed shared-mesh gradient buffer, taken
// directly over peer memory in ONE kernel: every rank's buffer is mapped into every process (symmetric memory over NVLink /
// NVSwitch).  Rank r owns slice r of the message:
//   peer path       float4 loads of slice r from every rank's buffer, summed in rank order (so every rank ends up with the
//                   same bits), float4 stores of the sums into slice r of every buffer;
//   multicast path  (a multicast mapping of the buffers is given) one multimem.ld_reduce per float4 -- the NVSwitch adds
//                   the copies on the way -- and one multimem.st, which the switch delivers to every copy.
// Two traversals of NVLink for 1/world of the message per rank, against the channel protocol of a general-purpose
// collective: at the 16 MB of config c5 the NCCL all-reduce is latency-bound (0.095 ms at 8 GPUs for a 0.49 ms step).
//
// Handshakes (flag words that only ever grow -- epoch numbers -- so nothing is reset and a fast peer that is already in
// its next call cannot be mistaken for a slow one):
//   before the sums   "my backward has finished": CTA 0, one thread per peer, raises this rank's flag in the peer's pad
//                     (release, system scope) and waits for the peer's flag in this rank's pad (acquire); then it opens
//                     the gate word the other CTAs of this kernel poll.  All CTAs are resident at once (grid <= 4 per SM).
//   after the sums    the CTA that finishes last (ticket) runs the same handshake on a second set of flags -- "every rank
//                     has stored its slice everywhere" -- and advances the epoch.
// A peer that never arrives must not hang the GPU for good: a wait gives up after ~30 s and records whom it waited for.
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../include/dirt_b200.h"

namespace dirt {

int set_error(int status, const char *what, cudaError_t e);      // capi.cu
void note_launches(uint64_t n);

namespace {

constexpr int kMaxWorld = 32;
constexpr long long kGiveUpCycles = 60000000000ll;      // ~30 s at 2 GHz
enum { kEpochCalls = 0, kEpochError = 1, kEpochGate = 2, kEpochDone = 3 };

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// one thread per peer (threadIdx.x < world); phase 0 / 1 use flag words [phase * world + sender] of every pad
__device__ __forceinline__ void handshake(uint32_t *const *signals, uint32_t *epoch, int rank, int world, int phase, uint32_t e,
                                          long long give_up)
{
    const int p = threadIdx.x;
    if (p < world) {
        __threadfence_system();                                   // what this rank wrote before is visible first
        st_release_sys(signals[p] + phase * world + rank, e);
        const uint32_t *mine = signals[rank] + phase * world + p;
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(mine) - e) < 0) {
            if (clock64() - t0 > give_up) { atomicMax(epoch + kEpochError, (uint32_t)(p + 1)); break; }
        }
    }
}

template <bool MULTICAST>
__global__ void __launch_bounds__(256) k_peer_allreduce(float *const *buffers, float *multicast, uint32_t *const *signals, uint32_t *epoch,
                                                        int rank, int world, long long count4, long long give_up)
{
    const uint32_t e = epoch[kEpochCalls] + 1u;                   // written only by the last CTA of the previous call
    if (blockIdx.x == 0) {
        handshake(signals, epoch, rank, world, 0, e, give_up);
        __syncthreads();
        if (threadIdx.x == 0) st_release_gpu(epoch + kEpochGate, e);
    } else {
        if (threadIdx.x == 0) {
            const long long t0 = clock64();
            while ((int32_t)(ld_acquire_gpu(epoch + kEpochGate) - e) < 0)
                if (clock64() - t0 > 2 * give_up) break;
        }
        __syncthreads();
    }
    // slice of this rank, in float4 units
    const long long per = (count4 + world - 1) / world;
    const long long lo = per * rank, hi = lo + per < count4 ? lo + per : count4;
    for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (long long)gridDim.x * blockDim.x) {
        if (MULTICAST) {
            float4 v;
            float4 *mc = reinterpret_cast<float4 *>(multicast) + i;
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                         :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
        } else {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int p = 0; p < world; ++p) {                     // fixed order: every rank computes the same bits
                const float4 v = __ldcg(reinterpret_cast<const float4 *>(buffers[p]) + i);
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
            for (int p = 0; p < world; ++p) __stcg(reinterpret_cast<float4 *>(buffers[p]) + i, acc);
        }
    }
    // the CTA that finishes last tells the peers and advances the epoch
    __shared__ bool s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        s_last = atomicAdd(epoch + kEpochDone, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        handshake(signals, epoch, rank, world, 1, e, give_up);
        __syncthreads();
        if (threadIdx.x == 0) { epoch[kEpochDone] = 0u; epoch[kEpochCalls] = e; }
    }
}

}  // namespace
}  // namespace dirt

extern "C" int dirt_allreduce_peer(const DirtPeerAllreduceArgs *a, void *stream)
{
    if (!a) return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "null arguments", cudaSuccess);
    if (a->world < 1 || a->world > dirt::kMaxWorld || a->rank < 0 || a->rank >= a->world)
        return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "allreduce_peer: 1 <= world <= 32 and 0 <= rank < world", cudaSuccess);
    if (a->count < 0 || (a->count & 3)) return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "allreduce_peer: count must be a multiple of 4 floats", cudaSuccess);
    if (a->world == 1 || a->count == 0) return DIRT_STATUS_OK;
    if (!a->buffers || !a->signals || !a->epoch) return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "allreduce_peer: null pointer", cudaSuccess);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long count4 = a->count / 4;
    int sms = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long g = ((count4 + a->world - 1) / a->world + 255) / 256;
    if (g > (long long)sms * 4) g = (long long)sms * 4;           // all CTAs resident at once: the gate below relies on it
    if (g < 1) g = 1;
    // DIRT_B200_PEER_TIMEOUT_MS: how long a handshake waits for a peer before it gives up (default ~30 s)
    long long give_up = dirt::kGiveUpCycles;
    if (const char *env = std::getenv("DIRT_B200_PEER_TIMEOUT_MS")) {
        const long long ms = std::atoll(env);
        if (ms > 0) give_up = ms * 2000000ll;
    }
    if (a->multicast)
        dirt::k_peer_allreduce<true><<<(unsigned)g, 256, 0, st>>>(a->buffers, a->multicast, a->signals, a->epoch, a->rank, a->world, count4, give_up);
    else
        dirt::k_peer_allreduce<false><<<(unsigned)g, 256, 0, st>>>(a->buffers, nullptr, a->signals, a->epoch, a->rank, a->world, count4, give_up);
    dirt::note_launches(1);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? DIRT_STATUS_OK : dirt::set_error(DIRT_STATUS_CUDA_ERROR, "allreduce_peer launch", e);
}
