// usv_fused.cuh -- one VecEnv.step() of 'usv-asmc-simple' as ONE persistent, warp-specialised kernel.
//
// The split path (usv_kernels.cuh) runs dynamics_kernel (one thread per env: a long dependent float64 chain, FP64-pipe /
// latency bound, 120 registers) and then sensor_kernel (one warp per env: short independent tests, instruction-issue
// bound, <= 80 registers) back to back; neither fills the SM on its own and the hand-over crosses L2.  Here both run
// at the same time on every SM, one CTA per SM:
//
//   dynamics warps  (FU_D warps, the HIGHEST warp ids: the sub-partition arbiter prefers them, and their dependent
//                   chain is what the step's latency hangs on).  setmaxnreg.inc gives them FU_DYN_REGS registers.  Each
//                   takes batches of 32 consecutive envs (thread per env, coalesced SoA columns), runs the 20 ASMC
//                   sub-steps + wrapper_tail (simple_env_asmc.py:18-27, simple_env.py:310-346), stores the state
//                   columns, writes the hand-over (pose, heading cos / sin, partial reward, flags, 15 header columns)
//                   into a slot of a shared-memory ring and arrives on the slot's `full` mbarrier.
//   sensor warps    (FU_S warps, setmaxnreg.dec to FU_SENS_REGS).  Each draws env tickets from a shared-memory counter
//                   (one env AHEAD: the env's obstacle row is already in flight with cp.async while the current env is
//                   processed), waits for the env's batch (`full`), copies the hand-over into registers, arrives on
//                   the slot's `empty` mbarrier (32 arrivals free the slot for the dynamics warp that needs it
//                   next), then runs the range-sensor pass (usv_sensor.cuh) and finish_env (flags, reward, episode
//                   statistics, observation row, auto-reset).
//
// All waits are inside one CTA (every warp of the CTA is resident), so there is no inter-CTA dependency.  The
// mbarrier waits are bounded: a wait that does not complete within ~1 s traps (a hang would cost a GPU lease).
// Batches: CTA c of G owns batches c, c + G, c + 2G, ... of the launch's env range [lo, hi); a 65 536-env shard is 2048
// batches = 13.8 per SM.
#pragma once
#include "usv_kernels.cuh"

namespace usv {

#ifndef USV_FUSED_DYN_WARPS
#define USV_FUSED_DYN_WARPS 4
#endif
#ifndef USV_FUSED_SENS_WARPS
#define USV_FUSED_SENS_WARPS 20
#endif
#ifndef USV_FUSED_DYN_REGS
#define USV_FUSED_DYN_REGS 120
#endif
#ifndef USV_FUSED_SENS_REGS
#define USV_FUSED_SENS_REGS 72
#endif
constexpr int FU_D = USV_FUSED_DYN_WARPS;    // one or two warpgroups (setmaxnreg is a warpgroup-wide instruction)
constexpr int FU_S = USV_FUSED_SENS_WARPS;
constexpr int FU_WARPS = FU_D + FU_S;
constexpr int FU_THREADS = FU_WARPS * 32;
constexpr int FU_RING = 2 * FU_D;            // hand-over slots: every dynamics warp can run one batch ahead of the sensors
static_assert(FU_D % 4 == 0 && FU_S % 4 == 0, "warp roles are whole warpgroups");
static_assert((FU_RING & (FU_RING - 1)) == 0, "ring size is a power of two");
static_assert(FU_D * USV_FUSED_DYN_REGS + FU_S * USV_FUSED_SENS_REGS <= 2048, "64 K registers per SM");

struct alignas(16) HandBatch {  // hand-over of one batch of 32 envs: structure of arrays (the dynamics lanes store conflict-free)
    double x[32], y[32], cs[32], sn[32], rew[32];
    unsigned flags[32];
    float head[32][HDR_PAD];    // observation columns 0..14 (simple_env.py:91-96), one 64-byte row per env
};

template <typename ObT>
struct alignas(16) FusedStage {  // what a sensor warp prefetches for its next env
    ObT obst[OBST_ROW];          // x[32] y[32] radius[32]
    int nob;                     // + the 16-byte chunk of the sensor-input row that holds n_obs
    int pad[3];
};

template <typename ObT>
struct alignas(128) FusedSmem {
    double2 raytab[N_SENS];
    SensorScratch<float> scratch[FU_WARPS];
    FusedStage<ObT> stage[FU_WARPS];
    HandBatch ring[FU_RING];
    unsigned long long full[FU_RING];   // mbarrier, 1 arrival: the batch's hand-over is in the slot
    unsigned long long empty[FU_RING];  // mbarrier, 32 arrivals: every env of the batch has been picked up
    int ticket;
};

// ---- mbarrier (shared::cta), generic proxy only ------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {  // .release.cta
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {  // .acquire.cta; the hardware suspends the warp
    const uint32_t a = smem_addr(bar);
    for (int spin = 0;; ++spin) {
        uint32_t done;
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
            : "=r"(done)
            : "r"(a), "r"(parity)
            : "memory");
        if (done) return;
        if (spin > (1 << 22)) __trap();  // never in a correct run: fail loudly instead of hanging the GPU
    }
}

template <typename ObT>
__device__ __forceinline__ void fused_prefetch(FusedStage<ObT>& st, const StatePtrs& s, long long i, int lane) {
    constexpr int TABLE_CHUNKS = (int)(sizeof(ObT) * OBST_ROW / 16);  // 24 (float32 table) or 48 (float64)
    constexpr int TAIL_CHUNK = TABLE_CHUNKS + (int)((sizeof(Hand) + HDR_PAD * sizeof(float)) / 16);  // n_obs | target | pad
    static_assert(sizeof(FusedStage<ObT>) == 16 * (TABLE_CHUNKS + 1), "stage = table + one chunk");
    const char* row = reinterpret_cast<const char*>(obst_row<ObT>(s, i));
    char* dst = reinterpret_cast<char*>(&st);
#pragma unroll
    for (int c = lane; c < TABLE_CHUNKS + 1; c += 32)
        __pipeline_memcpy_async(dst + 16 * c, row + 16 * (c < TABLE_CHUNKS ? c : TAIL_CHUNK), 16);
    __pipeline_commit();
}

template <bool OBST, typename ObT>
__global__ void __launch_bounds__(FU_THREADS, 1)
step_fused_kernel(StatePtrs s, StepFlags fl, const float* __restrict__ actions, float* __restrict__ obs,
                  float* __restrict__ reward, uint8_t* __restrict__ terminated, uint8_t* __restrict__ truncated,
                  float* __restrict__ terminal_obs) {
    extern __shared__ __align__(128) unsigned char fused_smem_raw[];
    FusedSmem<ObT>& sm = *reinterpret_cast<FusedSmem<ObT>*>(fused_smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long n_batches = (fl.hi - fl.lo + 31) >> 5;
    // batches of this CTA: blockIdx.x, blockIdx.x + gridDim.x, ...
    const int nb = (n_batches > (long long)blockIdx.x)
                       ? (int)((n_batches - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;

    if (OBST && tid < N_SENS) sm.raytab[tid] = s.raytab[tid];
    if (OBST && lane == 0) sm.scratch[warp].hpair[PAIR_SLOTS - 1] = Reading<float>::no_hit();  // the "no hit" slot
    if (tid == 0) {
        for (int r = 0; r < FU_RING; ++r) {
            mbar_init(&sm.full[r], 1);
            mbar_init(&sm.empty[r], 32);
        }
        sm.ticket = 0;
    }
    __syncthreads();

    if (warp >= FU_S) {
        // =================================== dynamics role ===================================================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(USV_FUSED_DYN_REGS));
        for (int q = warp - FU_S; q < nb; q += FU_D) {
            const long long i = fl.lo + (((long long)blockIdx.x + (long long)q * gridDim.x) << 5) + lane;
            const bool valid = i < fl.hi;
            Hand hd;
            float4 hrow[4];
            if (valid) dyn_env<false>(s, fl, actions, i, hd, hrow);
            const int slot = q & (FU_RING - 1);
            if (q >= FU_RING) mbar_wait(&sm.empty[slot], (unsigned)((q / FU_RING) - 1) & 1u);
            if (valid) {
                HandBatch& hb = sm.ring[slot];
                hb.x[lane] = hd.x;
                hb.y[lane] = hd.y;
                hb.cs[lane] = hd.cs;
                hb.sn[lane] = hd.sn;
                hb.rew[lane] = hd.rew_part;
                hb.flags[lane] = hd.flags;
                float4* dst = reinterpret_cast<float4*>(hb.head[lane]);
#pragma unroll
                for (int k = 0; k < 4; ++k) dst[k] = hrow[k];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.full[slot]);  // release: the lanes' state columns and hand-over are visible
        }
        return;
    }

    // ======================================= sensor role =========================================================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(USV_FUSED_SENS_REGS));
    constexpr unsigned FULL = 0xFFFFFFFFu;
    SensorScratch<float>& sc = sm.scratch[warp];
    FusedStage<ObT>& st = sm.stage[warp];
    const int n_local = nb * 32;  // env tickets of this CTA: ticket t = env (t & 31) of the CTA's batch (t >> 5)
    auto draw = [&]() -> int {
        int t = 0;
        if (lane == 0) t = atomicAdd(&sm.ticket, 1);
        return __shfl_sync(FULL, t, 0);
    };
    auto env_of = [&](int t) -> long long {
        return fl.lo + (((long long)blockIdx.x + (long long)(t >> 5) * gridDim.x) << 5) + (t & 31);
    };
    int t = draw();
    if (OBST && t < n_local) {
        const long long i0 = env_of(t);
        if (i0 < fl.hi) fused_prefetch<ObT>(st, s, i0, lane);
    }
    while (t < n_local) {
        const int t_next = draw();
        const int q = t >> 5, b = t & 31;
        const long long i = env_of(t);
        const bool valid = i < fl.hi;  // the launch's last batch may be ragged
        int nob = 0;
        double ox = 0.0, oy = 0.0, rr = 0.0;
        if (OBST) {
            // ---- this env's obstacle row: stage -> registers, then the stage takes the next env's row ----------
            __pipeline_wait_prior(0);
            __syncwarp();
            if (valid) {
                nob = st.nob;
                ox = (double)st.obst[lane];
                oy = (double)st.obst[MAX_OBST + lane];
                rr = (double)st.obst[2 * MAX_OBST + lane];
            }
            __syncwarp();
            if (t_next < n_local) {
                const long long i1 = env_of(t_next);
                if (i1 < fl.hi) fused_prefetch<ObT>(st, s, i1, lane);  // in flight during this env's pass
            }
        }
        // ---- the dynamics warp's hand-over ----------------------------------------------------------------
        const int slot = q & (FU_RING - 1);
        mbar_wait(&sm.full[slot], (unsigned)(q / FU_RING) & 1u);
        const HandBatch& hb = sm.ring[slot];
        const double x = hb.x[b], y = hb.y[b], cs = hb.cs[b], sn = hb.sn[b], rew_part = hb.rew[b];
        const unsigned flags = hb.flags[b];
        const float head = hb.head[b][lane & (HDR_PAD - 1)];
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[slot]);
        if (valid) {
            float sens[4] = {1.0f, 1.0f, 1.0f, 1.0f};  // max range / 100 (simple_env.py:83); ignore_obstacles: all rays (:222-224)
            bool term = false, collided = false;
            if (OBST) sensor_warp<float>(sc, sm.raytab, lane, nob, x, y, cs, sn, ox, oy, rr, sens, 0.05, term, collided);
            finish_env<ObT>(s, fl, i, lane, head, sens, term, collided, rew_part, flags, obs, reward, terminated, truncated,
                            terminal_obs);
        }
        t = t_next;
    }
}

}  // namespace usv
