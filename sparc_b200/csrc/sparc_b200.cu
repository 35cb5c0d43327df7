// sparc_b200.cu — C-ABI (include/sparc_b200.h) over the fused sm_100a step kernel.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -shared
//        -Xcompiler -fPIC  (see sparc_b200/build.py).  -fmad=false keeps the fp64 scalar path
// free of contraction so it rounds like the reference's Python floats; the fp32 stencil uses
// explicit fmaf().
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/sparc_b200.h"
#include "state_layout.h"
#include "step_common.cuh"
#include "step_variants.h"
#include <atomic>

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char *fmt, const char *detail = "")
{
    snprintf(g_err, sizeof(g_err), fmt, detail);
    return code;
}

int cuda_fail(cudaError_t e, const char *what)
{
    snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    return SPARC_E_CUDA;
}

#define CUDA_TRY(expr)                                      \
    do {                                                    \
        cudaError_t _e = (expr);                            \
        if (_e != cudaSuccess) return cuda_fail(_e, #expr); \
    } while (0)

struct FieldDesc {
    const char *name;
    char dtype;
    size_t elem;
};

const FieldDesc kFields[] = {
#define X(name, ctype, dt, init) {#name, dt, sizeof(ctype)},
    SPARC_STATE_FIELDS(X)
#undef X
#define X(name, ctype, dt) {"prm_" #name, dt, sizeof(ctype)},
    SPARC_ENV_PARAM_FIELDS(X)
#undef X
    {"episode", 'i', sizeof(int32_t)},
    {"T", 'f', sizeof(float)},
    {"mech_traj", 'd', sizeof(double)},
    {"mech_trajv", 'd', sizeof(double)},
    {"mech_ckpt", 'd', sizeof(double)},
    {"mech_prog", 'i', sizeof(int32_t)},
};
constexpr int kNumFields = (int)(sizeof(kFields) / sizeof(kFields[0]));
constexpr int kFieldT = kNumFields - 5, kFieldTraj = kNumFields - 4, kFieldTrajV = kNumFields - 3, kFieldCkpt = kNumFields - 2;

constexpr size_t kAlign = 256;
constexpr int SPARC_SELF_MECH_US = 8;
constexpr int SPARC_ZERO_COPY_ENVS = 32;

int padded_segments(int n_segments) { return (int)sparc_align_up((size_t)n_segments, 32); }

size_t field_count(int i, int n_envs, int n_segments)
{
    if (i == kFieldT) return (size_t)n_envs * padded_segments(n_segments);
    if (i == kFieldTraj || i == kFieldTrajV) return (size_t)n_envs * SPARC_MECH_EPOCH;
    if (i == kFieldCkpt) return (size_t)n_envs * SPARC_MECH_NCKPT * 2;
    return (size_t)n_envs;
}

// offset of field idx; idx == kNumFields gives total bytes
size_t field_offset(int idx, int n_envs, int n_segments)
{
    size_t off = 0;
    for (int i = 0; i < idx; i++) off = sparc_align_up(off + field_count(i, n_envs, n_segments) * kFields[i].elem, kAlign);
    return off;
}

SparcStatePtrs make_ptrs(void *state, int n_envs, int n_segments)
{
    SparcStatePtrs S;
    char *base = (char *)state;
    int idx = 0;
#define X(name, ctype, dt, init) S.name = (ctype *)(base + field_offset(idx++, n_envs, n_segments));
    SPARC_STATE_FIELDS(X)
#undef X
#define X(name, ctype, dt) S.prm_##name = (ctype *)(base + field_offset(idx++, n_envs, n_segments));
    SPARC_ENV_PARAM_FIELDS(X)
#undef X
    S.episode = (int32_t *)(base + field_offset(idx++, n_envs, n_segments));
    S.T = (float *)(base + field_offset(idx++, n_envs, n_segments));
    S.mech_traj = (double *)(base + field_offset(idx++, n_envs, n_segments));
    S.mech_trajv = (double *)(base + field_offset(idx++, n_envs, n_segments));
    S.mech_ckpt = (double *)(base + field_offset(idx++, n_envs, n_segments));
    S.mech_prog = (int32_t *)(base + field_offset(idx++, n_envs, n_segments));
    S.n_pad = padded_segments(n_segments);
    return S;
}

// ---- small kernels --------------------------------------------------------------------------
__global__ void reset_kernel(const __grid_constant__ SparcParams P, SparcStatePtrs S, int n_envs,
                             const uint8_t *mask, int first_time)
{
    int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_envs) return;
    if (mask && !mask[e]) return;
    if (first_time == 3) {
        // The REFERENCE's reset() (wire_edm.py:106-114): a new EDMState with workpiece_position = initial_gap; what the
        // module objects keep on themselves survives: short-circuit countdowns and the current cache
        // (ignition.py:75-83), prev_accel (mechanics.py:60), debris volume / flow condition / ion channel / hysteresis
        // memory (dielectric.py:69-80), convection coefficients, flow memory and zone-mean cache (wire.py:204-225).
        S.wp[e] = S.prm_initial_gap[e]; S.wire[e] = 0.0; S.v[e] = 0.0;
        S.voltage[e] = 0.0; S.current[e] = 0.0;
        S.target_voltage[e] = 0.0; S.on_time[e] = 0.0; S.off_time[e] = 0.0; S.target_delta[e] = 0.0;
        S.unwind_v[e] = 0.2;
        S.debris_density[e] = 0.0; S.cavity_volume[e] = 0.0; S.last_crater_volume[e] = 0.0;
        S.spark_y[e] = -1.0; S.wire_avg_T[e] = 0.0;
        S.stat_crater_sum[e] = 0.0; S.win_vsum[e] = 0.0;
        S.time[e] = 0; S.t_servo[e] = 0; S.t_open_v[e] = 0; S.t_spark_ign[e] = 0; S.t_spark_end[e] = 0;
        S.spark_state[e] = 0; S.spark_dur[e] = 0; S.crit_count[e] = 0; S.current_mode[e] = 0; S.flags[e] = 0;
        S.stat_sparks[e] = 0; S.stat_short_us[e] = 0; S.stat_short_events[e] = 0; S.stat_pulse_short[e] = 0;
        S.stat_steps[e] = 0; S.win_sparks[e] = 0; S.win_short_us[e] = 0; S.win_steps[e] = 0;
        S.episode[e] = S.episode[e] + 1;
        S.mech_prog[e] = 0;
        return;
    }
    if (first_time == 2) {                      // sparc_init: per-env parameters start as the engine's SparcParams
#define X(name, ctype, dt) S.prm_##name[e] = (ctype)P.name;
        SPARC_ENV_PARAM_FIELDS(X)
#undef X
    }
    const double env_initial_gap = S.prm_initial_gap[e];
#define X(name, ctype, dt, init) S.name[e] = (ctype)(init);
    SPARC_STATE_FIELDS(X)
#undef X
    S.episode[e] = first_time ? 0 : S.episode[e] + 1;
    S.mech_prog[e] = 0;
}

__global__ void reset_T_kernel(SparcStatePtrs S, int n_envs, float spool, const uint8_t *mask)
{
    size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t total = (size_t)n_envs * S.n_pad;
    if (idx >= total) return;
    int e = (int)(idx / S.n_pad);
    if (mask && !mask[e]) return;
    S.T[idx] = spool;
}

// one warp per env: max over the temperature row + scalar gather
__global__ void observe_kernel(SparcStatePtrs S, int n_envs, int n_segments, float *obs)
{
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (warp >= n_envs) return;
    const float *row = S.T + (size_t)warp * S.n_pad;
    float m = -3.0e38f;
    for (int i = lane; i < n_segments; i += 32) m = fmaxf(m, row[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) {
        int e = warp;
        int steps = S.win_steps[e];
        obs[0 * (size_t)n_envs + e] = (float)(S.wp[e] - S.wire[e]);
        obs[1 * (size_t)n_envs + e] = (float)S.v[e];
        obs[2 * (size_t)n_envs + e] = steps > 0 ? (float)(S.win_vsum[e] / (double)steps) : 0.0f;
        obs[3 * (size_t)n_envs + e] = (float)S.win_sparks[e];
        obs[4 * (size_t)n_envs + e] = (float)S.win_short_us[e];
        obs[5 * (size_t)n_envs + e] = m;
        obs[6 * (size_t)n_envs + e] = (float)S.debris_density[e];
        obs[7 * (size_t)n_envs + e] = (float)S.flags[e];
        obs[8 * (size_t)n_envs + e] = __int_as_float(S.time[e]);
        obs[9 * (size_t)n_envs + e] = __int_as_float(S.spark_state[e]);
        obs[10 * (size_t)n_envs + e] = __int_as_float(S.t_servo[e]);
        obs[11 * (size_t)n_envs + e] = 0.0f;
    }
}

__global__ void reduce_stats_kernel(SparcStatePtrs S, int n_envs, double *out)
{
    double acc[SPARC_N_STATS] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n_envs; e += gridDim.x * blockDim.x) {
        int fl = S.flags[e];
        acc[0] += S.stat_sparks[e];
        acc[1] += S.stat_short_us[e];
        acc[2] += S.stat_short_events[e];
        acc[3] += S.stat_crater_sum[e];
        acc[4] += S.stat_pulse_short[e];
        acc[5] += S.stat_steps[e];
        acc[6] += (fl & SPARC_F_BROKEN) ? 1.0 : 0.0;
        acc[7] += (fl & SPARC_F_TARGET) ? 1.0 : 0.0;
    }
#pragma unroll
    for (int k = 0; k < SPARC_N_STATS; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
        if ((threadIdx.x & 31) == 0) atomicAdd(&out[k], acc[k]);
    }
}

// FP32 issue-rate probe: 8 independent FFMA chains per thread (roofline denominator)
__global__ void fma_peak_kernel(float *out, float a, float b, int iters)
{
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = threadIdx.x * 1e-3f + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) x[i] = fmaf(x[i], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---- step kernel dispatch -------------------------------------------------------------------
using Variant = sparc::StepVariant;

// Region layout of the wire in the 64*W halves of an environment's lanes (SparcLayout, step_common.cuh).
// A region of n cells takes nh >= ceil(n/H) halves, nh*H - n of them short (one pad each).
// ALIGNED (three regions: below the zone, the zone, above it; per-half coefficients): pads only in y
// halves (odd h) and the wire ends on a y half, so the stencil needs no x-side pad handling.
// Otherwise: ONE region, the first p halves short, per-cell coefficients (the general kernel variant).
int odd_halves(int hb, int nh) { return (hb + nh) / 2 - hb / 2; }    // odd h in [hb, hb + nh)

bool try_aligned(const int n[3], const int nh[3], int H, int NH, sparc::SparcLayout *L, const int cb[3])
{
    int hb = 0;
    L->hb[0] = 0;
    for (int r = 0; r < 3; r++) {
        const int p = nh[r] * H - n[r];
        if (p < 0 || p > odd_halves(hb, nh[r])) return false;
        L->p[r] = p;
        L->cb[r] = cb[r];
        hb += nh[r];
        L->hb[r + 1] = hb;
    }
    return hb <= NH && hb % 2 == 0;
}

bool compute_layout(const SparcParams *p, int K, int W, sparc::SparcLayout *L)
{
    const int H = K / 2, NH = 64 * W, N = p->n_segments;
    int zs = p->actual_zone_start, ze = p->actual_zone_end;
    if (!(ze > zs)) { zs = 1; ze = 1; }          // empty zone
    zs = zs < 1 ? 1 : (zs > N ? N : zs);
    ze = ze < zs ? zs : (ze > N ? N : ze);
    const int cb[4] = {1, zs, ze, N};
    int n[3], nh[3];
    for (int r = 0; r < 3; r++) { n[r] = cb[r + 1] - cb[r]; nh[r] = (n[r] + H - 1) / H; }
    L->aligned = 1;
    if (try_aligned(n, nh, H, NH, L, cb)) return true;
    for (int r = 2; r >= 0; r--) {               // one more half in one region: fixes an odd total or too few y halves
        if (n[r] == 0) continue;
        int nh2[3] = {nh[0], nh[1], nh[2]};
        nh2[r] += 1;
        if (try_aligned(n, nh2, H, NH, L, cb)) return true;
    }
    // general layout
    const int M = N - 1;
    int nh0 = (M + H - 1) / H, p0 = nh0 * H - M;
    if (p0 > nh0 || nh0 > NH) return false;
    L->aligned = 0;
    L->hb[0] = 0; L->hb[1] = L->hb[2] = L->hb[3] = nh0;
    L->p[0] = p0; L->p[1] = L->p[2] = 0;
    L->cb[0] = 1; L->cb[1] = L->cb[2] = N;
    return true;
}

// fewest warps per env, then fewest cells per lane, whose 32*W*K slots hold the n_segments-1 dynamic
// cells in a valid layout (aligned if the slack allows, else one region with per-cell coefficients)
const Variant *pick_variant(const SparcParams *p, sparc::SparcLayout *L_out)
{
    const int M = p->n_segments - 1;
    const Variant *(*tables[4])(int *) = {sparc::step_variants_w1, sparc::step_variants_w2, sparc::step_variants_w4,
                                          sparc::step_variants_w8};
    const Variant *best = nullptr;
    for (int t = 0; t < 4; t++) {
        int n = 0;
        const Variant *tab = tables[t](&n);
        for (int i = 0; i < n; i++) {
            const Variant &v = tab[i];
            if (v.K * 32 * v.W < M) continue;
            if (best && !(v.W < best->W || (v.W == best->W && v.K < best->K))) continue;
            sparc::SparcLayout L;
            if (!compute_layout(p, v.K, v.W, &L)) continue;
            best = &v;
            if (L_out) *L_out = L;
        }
    }
    return best;
}

// Grid of the step kernel: one warp (W == 1) or block (W > 1) per environment.
void launch_geometry(const Variant *v, int n_envs, int *blocks, int *threads)
{
    if (v->W == 1) {
        const int wpb = SPARC_BLOCK_THREADS / 32;
        *threads = SPARC_BLOCK_THREADS;
        *blocks = (n_envs + wpb - 1) / wpb;
    } else {
        *threads = 32 * v->W;
        *blocks = n_envs;
    }
}

int check_common(const void *state, const SparcParams *p, int n_envs)
{
    if (!state || !p) return fail(SPARC_E_INVALID, "null state/params pointer");
    if (n_envs <= 0) return fail(SPARC_E_INVALID, "n_envs must be positive");
    if (p->n_segments < 2) return fail(SPARC_E_UNSUPPORTED, "n_segments < 2 is not supported");
    return SPARC_OK;
}

// Tag of one (pre-pass, step) launch pair: progress values are tag + blocks published, so a value
// left behind by an earlier launch can never be mistaken for progress of this one.
uint32_t next_mech_seq()
{
    static std::atomic<uint32_t> counter{0};
    uint32_t s;
    do { s = (counter.fetch_add(1) + 1) & 0x3ffffffu; } while (s == 0);
    return s << 6;
}

int launch_step(void *state, const SparcParams *p, const SparcActions *actions, int n_envs, int n_us,
                uint64_t seed, uint32_t env_id_base, const double *draws, void *stream, float *tlog = nullptr,
                int log_every = 0, int n_log_envs = 0, double *signals = nullptr, int sample_every = 0,
                float *obs = nullptr, const SparcController *ctrl = nullptr)
{
    int rc = check_common(state, p, n_envs);
    if (rc) return rc;
    if (!actions || !actions->servo || !actions->target_voltage || !actions->current_mode ||
        !actions->on_time || !actions->off_time)
        return fail(SPARC_E_INVALID, "null action array");
    if (n_us < 0) return fail(SPARC_E_INVALID, "n_us must be >= 0");
    if (n_us == 0) return SPARC_OK;
    if (ctrl && ctrl->kind != 0) {
        if (ctrl->kind != 1) return fail(SPARC_E_INVALID, "unknown controller kind");
        if (n_us > SPARC_MECH_EPOCH) return fail(SPARC_E_UNSUPPORTED, "in-kernel controller: single-launch calls only (n_us <= 1024)");
    }
    sparc::SparcLayout lay;
    const Variant *v = pick_variant(p, &lay);
    if (!v) return fail(SPARC_E_UNSUPPORTED, "n_segments too large (max 6657 wire segments)");
    SparcStatePtrs S = make_ptrs(state, n_envs, p->n_segments);
    const sparc::StepFn fn = v->fn[(tlog || signals) ? 1 : 0][draws ? 1 : 0][lay.aligned ? 1 : 0];
    int blocks, threads;
    launch_geometry(v, n_envs, &blocks, &threads);
    // one (mechanics pre-pass, fused step) pair per chunk of at most SPARC_MECH_EPOCH µs
    int t = 0;
    while (t < n_us) {
        int n = (n_us - t < SPARC_MECH_EPOCH) ? n_us - t : SPARC_MECH_EPOCH;
        sparc::StepArgs a;
        a.act = *actions;
        a.n_envs = n_envs;
        a.n_us = n;
        a.seed = seed;
        a.env_id_base = env_id_base;
        a.draws = draws;
        a.tlog = tlog;
        a.log_every = log_every;
        a.n_log_envs = n_log_envs;
        a.it_base = t;
        a.signals = signals;              // rows are written by the step kernel itself (LOG instantiations)
        a.sample_every = signals ? sample_every : 0;
        a.total_us = n_us;
        a.obs = (t + n >= n_us) ? obs : nullptr;    // the last launch of the call writes the observation
        // The pre-pass and the step kernel OVERLAP: mech_kernel triggers programmatic launch completion
        // in its first instruction, so the step kernel (launched with programmatic stream
        // serialization) starts as soon as every pre-pass block is resident, and consumes the
        // trajectory block by block behind per-env progress tags (mech_prog) instead of waiting
        // for the pre-pass grid to drain.
        // Launches of at most SPARC_SELF_MECH_US µs (the drop-in's 1-µs step(), per-µs sampling) skip the
        // pre-pass kernel: mech_seq = 0 tells the step kernel to evaluate the few mechanics steps itself.
        const bool prepass = n > SPARC_SELF_MECH_US;
        a.mech_seq = prepass ? next_mech_seq() : 0u;
        a.lay = lay;
        if (ctrl) a.ctrl = *ctrl; else { a.ctrl.kind = 0; a.ctrl.reserved = 0; a.ctrl.desired_gap = a.ctrl.gain = a.ctrl.limit = 0.0; }
        if (prepass) {
            sparc::mech_launch(*p, S, a, (cudaStream_t)stream);
            CUDA_TRY(cudaGetLastError());         // never launch the consumer behind a producer that did not start
        }
        {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)blocks);
            cfg.blockDim = dim3((unsigned)threads);
            // LOG kernels stage rows in dynamic shared memory: two buffers of (N rounded up to 4) floats per env
            size_t dyn = 0;
            if (tlog || signals) {
                const size_t n_stage = ((size_t)p->n_segments + 3) & ~(size_t)3;
                dyn = (size_t)SPARC_LOG_NBUF(v->W) * n_stage * sizeof(float) * (v->W == 1 ? SPARC_BLOCK_THREADS / 32 : 1);
                if (dyn > 40 * 1024) CUDA_TRY(cudaFuncSetAttribute((const void *)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
            }
#ifdef SPARC_EXPERIMENT_KNOBS   /* experiments only: cap the resident blocks per SM with unused dynamic shared memory */
            static const char *exp_dyn = getenv("SPARC_EXP_DYNSMEM");
            if (exp_dyn && dyn == 0) {
                dyn = (size_t)atoi(exp_dyn);
                if (dyn > 40 * 1024) CUDA_TRY(cudaFuncSetAttribute((const void *)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
            }
#endif
            cfg.dynamicSmemBytes = dyn;
            cfg.stream = (cudaStream_t)stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            // SPARC_NO_PDL=1 (diagnostics): plain stream order, the step kernel starts after the pre-pass
            static const bool no_pdl = getenv("SPARC_NO_PDL") != nullptr;
            cfg.numAttrs = (no_pdl || !prepass) ? 0 : 1;
            CUDA_TRY(cudaLaunchKernelEx(&cfg, fn, *p, S, a));
        }
        t += n;
    }
    CUDA_TRY(cudaGetLastError());
    return SPARC_OK;
}

// ---- allocation-free variant of the host-buffer step (sparc_step_host_static) ---------------------
// The action staging arrays and the observation summary that sparc_step_host keeps in caller-owned device
// scratch live here instead: a file-scope array every GPU that loads the library holds until exit.
// Size: the scratch of SPARC_STATIC_MAX_ENVS envs (5 action arrays + SPARC_N_OBS observation rows, each padded
// to 256 bytes) = 1 114 112 bytes.  Shared by ALL calls on a device: calls that overlap on the GPU race.
constexpr size_t static_scratch_bytes(size_t n_envs)
{
    return ((n_envs * 4 + kAlign - 1) / kAlign * kAlign) * 5 + (n_envs * 4 * SPARC_N_OBS + kAlign - 1) / kAlign * kAlign;
}
static __device__ __align__(256) unsigned char g_static_scratch[static_scratch_bytes(SPARC_STATIC_MAX_ENVS)];

// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is
// looked up once per device, with that device current, and the kernels are given that pointer.
int static_scratch_ptr(void **out)
{
    constexpr int kMaxDev = 64;
    static void *ptrs[kMaxDev] = {};
    static std::atomic<bool> have[kMaxDev];
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDev) return fail(SPARC_E_UNSUPPORTED, "device ordinal beyond 63");
    if (!have[dev].load(std::memory_order_acquire)) {
        void *ptr = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void **)&ptr, g_static_scratch));
        if (!ptr) return fail(SPARC_E_CUDA, "cudaGetSymbolAddress returned a null address");
        ptrs[dev] = ptr;                          // every thread resolves the same address: a benign double write
        have[dev].store(true, std::memory_order_release);
    }
    *out = ptrs[dev];
    return SPARC_OK;
}

}  // namespace

extern "C" {

int sparc_abi_version(void) { return SPARC_ABI_VERSION; }
const char *sparc_last_error(void) { return g_err; }
int sparc_num_fields(void) { return kNumFields; }
int sparc_padded_segments(int n_segments) { return padded_segments(n_segments); }

int sparc_field_info(int idx, int n_envs, int n_segments, const char **name, char *dtype,
                     size_t *byte_offset, size_t *count)
{
    if (idx < 0 || idx >= kNumFields) return fail(SPARC_E_INVALID, "field index out of range");
    if (n_envs <= 0 || n_segments <= 0) return fail(SPARC_E_INVALID, "bad n_envs/n_segments");
    if (name) *name = kFields[idx].name;
    if (dtype) *dtype = kFields[idx].dtype;
    if (byte_offset) *byte_offset = field_offset(idx, n_envs, n_segments);
    if (count) *count = field_count(idx, n_envs, n_segments);
    return SPARC_OK;
}

size_t sparc_state_bytes(int n_envs, int n_segments)
{
    if (n_envs <= 0 || n_segments <= 0) return 0;
    return field_offset(kNumFields, n_envs, n_segments);
}

static int do_reset(void *state, const SparcParams *p, int n_envs, const uint8_t *mask_dev, void *stream,
                    int first_time)
{
    int rc = check_common(state, p, n_envs);
    if (rc) return rc;
    SparcStatePtrs S = make_ptrs(state, n_envs, p->n_segments);
    cudaStream_t st = (cudaStream_t)stream;
    reset_kernel<<<(n_envs + 127) / 128, 128, 0, st>>>(*p, S, n_envs, mask_dev, first_time);
    size_t total = (size_t)n_envs * S.n_pad;
    reset_T_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(S, n_envs, (float)p->spool_T, mask_dev);
    CUDA_TRY(cudaGetLastError());
    return SPARC_OK;
}

int sparc_init(void *state, const SparcParams *p, int n_envs, void *stream)
{
    return do_reset(state, p, n_envs, nullptr, stream, 2);
}

int sparc_reset(void *state, const SparcParams *p, int n_envs, const uint8_t *mask_dev, void *stream)
{
    return do_reset(state, p, n_envs, mask_dev, stream, 0);
}

int sparc_reseed(void *state, const SparcParams *p, int n_envs, const uint8_t *mask_dev, void *stream)
{
    return do_reset(state, p, n_envs, mask_dev, stream, 1);
}

int sparc_reset_reference(void *state, const SparcParams *p, int n_envs, const uint8_t *mask_dev, void *stream)
{
    return do_reset(state, p, n_envs, mask_dev, stream, 3);
}

size_t sparc_params_bytes(void) { return sizeof(SparcParams); }

int sparc_step(void *state, const SparcParams *p, const SparcActions *actions, int n_envs, int n_us,
               uint64_t seed, uint32_t env_id_base, void *stream)
{
    return launch_step(state, p, actions, n_envs, n_us, seed, env_id_base, nullptr, stream);
}

int sparc_step_injected(void *state, const SparcParams *p, const SparcActions *actions, int n_envs,
                        int n_us, const double *draws_dev, void *stream)
{
    if (!draws_dev) return fail(SPARC_E_INVALID, "null draws pointer");
    return launch_step(state, p, actions, n_envs, n_us, 0, 0, draws_dev, stream);
}

int sparc_step_logged(void *state, const SparcParams *p, const SparcActions *actions, int n_envs, int n_us,
                      uint64_t seed, uint32_t env_id_base, const double *draws_dev, float *tlog_dev, int log_every,
                      int n_log_envs, void *stream)
{
    if (!tlog_dev) return fail(SPARC_E_INVALID, "null log pointer");
    if (log_every <= 0 || n_log_envs < 0 || n_log_envs > n_envs) return fail(SPARC_E_INVALID, "bad log_every / n_log_envs");
    return launch_step(state, p, actions, n_envs, n_us, seed, env_id_base, draws_dev, stream, tlog_dev, log_every,
                       n_log_envs);
}

int sparc_step_sampled(void *state, const SparcParams *p, const SparcActions *actions, int n_envs, int n_us,
                       uint64_t seed, uint32_t env_id_base, const double *draws_dev, float *tlog_dev, int log_every,
                       int n_log_envs, double *signals_dev, int sample_every, void *stream)
{
    if (!signals_dev || sample_every <= 0) return fail(SPARC_E_INVALID, "signal log needs a buffer and sample_every > 0");
    if (tlog_dev) {
        if (log_every <= 0) return fail(SPARC_E_INVALID, "temperature log needs log_every > 0");
        if (n_log_envs < 0 || n_log_envs > n_envs) return fail(SPARC_E_INVALID, "n_log_envs out of range");
    }
    return launch_step(state, p, actions, n_envs, n_us, seed, env_id_base, draws_dev, stream, tlog_dev,
                       tlog_dev ? log_every : 0, tlog_dev ? n_log_envs : 0, signals_dev, sample_every);
}

int sparc_step_observe(void *state, const SparcParams *p, const SparcActions *actions, int n_envs, int n_us,
                       uint64_t seed, uint32_t env_id_base, const double *draws_dev, float *obs_dev, void *stream)
{
    if (!obs_dev) return fail(SPARC_E_INVALID, "null observation buffer");
    int rc = launch_step(state, p, actions, n_envs, n_us, seed, env_id_base, draws_dev, stream, nullptr, 0, 0, nullptr, 0,
                         obs_dev);
    if (rc || n_us > 0) return rc;
    return sparc_observe(state, n_envs, p->n_segments, obs_dev, stream);
}

int sparc_step_ex(void *state, const SparcParams *p, const SparcActions *actions, const SparcController *ctrl,
                  int n_envs, int n_us, uint64_t seed, uint32_t env_id_base, const double *draws_dev, float *obs_dev,
                  void *stream)
{
    int rc = launch_step(state, p, actions, n_envs, n_us, seed, env_id_base, draws_dev, stream, nullptr, 0, 0, nullptr, 0,
                         obs_dev, ctrl);
    if (rc || n_us > 0 || !obs_dev) return rc;
    return sparc_observe(state, n_envs, p->n_segments, obs_dev, stream);
}

int sparc_observe(const void *state, int n_envs, int n_segments, float *obs_dev, void *stream)
{
    if (!state || !obs_dev || n_envs <= 0 || n_segments <= 0) return fail(SPARC_E_INVALID, "bad observe arguments");
    SparcStatePtrs S = make_ptrs((void *)state, n_envs, n_segments);
    int threads = 128;
    int blocks = (int)(((size_t)n_envs * 32 + threads - 1) / threads);
    observe_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(S, n_envs, n_segments, obs_dev);
    CUDA_TRY(cudaGetLastError());
    return SPARC_OK;
}

int sparc_reduce_stats(const void *state, int n_envs, int n_segments, double *out_dev8, void *stream)
{
    if (!state || !out_dev8 || n_envs <= 0 || n_segments <= 0) return fail(SPARC_E_INVALID, "bad reduce arguments");
    SparcStatePtrs S = make_ptrs((void *)state, n_envs, n_segments);
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaMemsetAsync(out_dev8, 0, sizeof(double) * SPARC_N_STATS, st));
    int blocks = (n_envs + 255) / 256;
    if (blocks > 148 * 4) blocks = 148 * 4;
    reduce_stats_kernel<<<blocks, 256, 0, st>>>(S, n_envs, out_dev8);
    CUDA_TRY(cudaGetLastError());
    return SPARC_OK;
}

size_t sparc_host_action_pitch(int n_envs)
{
    if (n_envs <= 0) return 0;
    return sparc_align_up((size_t)n_envs * 4, kAlign);
}

size_t sparc_host_scratch_bytes(int n_envs)
{
    if (n_envs <= 0) return 0;
    size_t n = sparc_align_up((size_t)n_envs * 4, kAlign);
    return n * 5 + sparc_align_up((size_t)n_envs * 4 * SPARC_N_OBS, kAlign);
}

int sparc_step_host(void *state, const SparcParams *p, const SparcActions *ah, int n_envs, int n_us,
                    uint64_t seed, uint32_t env_id_base, void *scratch_dev, float *obs_host, void *stream)
{
    int rc = check_common(state, p, n_envs);
    if (rc) return rc;
    if (!ah || !scratch_dev || !obs_host) return fail(SPARC_E_INVALID, "null host action / scratch / obs pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_envs <= SPARC_ZERO_COPY_ENVS && n_us > 0) {
        // A handful of envs (the drop-in's single env): the kernel reads the five action scalars straight
        // from the pinned host arrays and writes the observation row straight into pinned host memory
        // (unified addressing: a pinned host pointer is a valid device pointer).  Two copy calls and their
        // completion latency cost more than the whole 1-us launch; for thousands of envs the staging copies
        // win (measured in round 1: +17 us per call without them at 4096 envs).
        rc = launch_step(state, p, ah, n_envs, n_us, seed, env_id_base, nullptr, stream, nullptr, 0, 0, nullptr, 0, obs_host);
        if (rc) return rc;
        CUDA_TRY(cudaStreamSynchronize(st));
        return SPARC_OK;
    }
    size_t n = sparc_align_up((size_t)n_envs * 4, kAlign);
    char *d = (char *)scratch_dev;
    SparcActions ad;
    ad.servo = (const float *)(d + 0 * n);
    ad.target_voltage = (const float *)(d + 1 * n);
    ad.current_mode = (const int32_t *)(d + 2 * n);
    ad.on_time = (const float *)(d + 3 * n);
    ad.off_time = (const float *)(d + 4 * n);
    float *obs_dev = (float *)(d + 5 * n);
    size_t bytes = (size_t)n_envs * 4;
    const char *h0 = (const char *)ah->servo;
    const bool packed = (const char *)ah->target_voltage == h0 + n && (const char *)ah->current_mode == h0 + 2 * n &&
                        (const char *)ah->on_time == h0 + 3 * n && (const char *)ah->off_time == h0 + 4 * n;
    if (packed) {
        // the five host arrays sit in one buffer with the scratch pitch (sparc_host_action_pitch): one copy
        CUDA_TRY(cudaMemcpyAsync(d, h0, 4 * n + bytes, cudaMemcpyHostToDevice, st));
    } else {
        CUDA_TRY(cudaMemcpyAsync((void *)ad.servo, ah->servo, bytes, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync((void *)ad.target_voltage, ah->target_voltage, bytes, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync((void *)ad.current_mode, ah->current_mode, bytes, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync((void *)ad.on_time, ah->on_time, bytes, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync((void *)ad.off_time, ah->off_time, bytes, cudaMemcpyHostToDevice, st));
    }
    // the observation is written by the step kernel's epilogue (same values as sparc_observe)
    rc = launch_step(state, p, &ad, n_envs, n_us, seed, env_id_base, nullptr, stream, nullptr, 0, 0, nullptr, 0, obs_dev);
    if (rc) return rc;
    if (n_us == 0) {
        rc = sparc_observe(state, n_envs, p->n_segments, obs_dev, stream);
        if (rc) return rc;
    }
    CUDA_TRY(cudaMemcpyAsync(obs_host, obs_dev, bytes * SPARC_N_OBS, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return SPARC_OK;
}

int sparc_static_max_envs(void) { return SPARC_STATIC_MAX_ENVS; }

int sparc_step_host_static(void *state, const SparcParams *p, const SparcActions *ah, int n_envs, int n_us,
                           uint64_t seed, uint32_t env_id_base, float *obs_host, void *stream)
{
    int rc = check_common(state, p, n_envs);
    if (rc) return rc;
    if (!ah || !obs_host) return fail(SPARC_E_INVALID, "null host action / obs pointer");
    if (!ah->servo || !ah->target_voltage || !ah->current_mode || !ah->on_time || !ah->off_time)
        return fail(SPARC_E_INVALID, "null action array");
    if (n_envs > SPARC_STATIC_MAX_ENVS) {
        snprintf(g_err, sizeof(g_err), "sparc_step_host_static: n_envs %d exceeds the static scratch (at most %d envs); "
                 "use sparc_step_host with caller-owned scratch", n_envs, SPARC_STATIC_MAX_ENVS);
        return SPARC_E_UNSUPPORTED;
    }
    void *scratch = nullptr;
    rc = static_scratch_ptr(&scratch);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t n = sparc_align_up((size_t)n_envs * 4, kAlign), bytes = (size_t)n_envs * 4;
    char *d = (char *)scratch;
    SparcActions ad;
    ad.servo = (const float *)(d + 0 * n);
    ad.target_voltage = (const float *)(d + 1 * n);
    ad.current_mode = (const int32_t *)(d + 2 * n);
    ad.on_time = (const float *)(d + 3 * n);
    ad.off_time = (const float *)(d + 4 * n);
    float *obs_dev = (float *)(d + 5 * n);
    // nothing in the scratch has to start at zero: the five action arrays are overwritten here and the
    // kernel writes all SPARC_N_OBS rows of every env (terminated ones included) before they are read back
    CUDA_TRY(cudaMemcpyAsync((void *)ad.servo, ah->servo, bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync((void *)ad.target_voltage, ah->target_voltage, bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync((void *)ad.current_mode, ah->current_mode, bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync((void *)ad.on_time, ah->on_time, bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync((void *)ad.off_time, ah->off_time, bytes, cudaMemcpyHostToDevice, st));
    rc = launch_step(state, p, &ad, n_envs, n_us, seed, env_id_base, nullptr, stream, nullptr, 0, 0, nullptr, 0, obs_dev);
    if (rc) return rc;
    if (n_us == 0) {
        rc = sparc_observe(state, n_envs, p->n_segments, obs_dev, stream);
        if (rc) return rc;
    }
    CUDA_TRY(cudaMemcpyAsync(obs_host, obs_dev, bytes * SPARC_N_OBS, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return SPARC_OK;
}

int sparc_measure_fp32_peak(void *scratch_dev, size_t scratch_bytes, double *tflops_out)
{
    if (!scratch_dev || !tflops_out) return fail(SPARC_E_INVALID, "null scratch/out pointer");
    int dev = 0, sms = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int threads = 512, bpsm = 4, iters = 1 << 16;
    int blocks = sms * bpsm;
    if ((size_t)blocks * threads * sizeof(float) > scratch_bytes) return fail(SPARC_E_INVALID, "scratch too small");
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 4; rep++) {
        CUDA_TRY(cudaEventRecord(e0, 0));
        fma_peak_kernel<<<blocks, threads>>>((float *)scratch_dev, 1.0001f, 0.5f, iters);
        CUDA_TRY(cudaEventRecord(e1, 0));
        CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        double tf = 2.0 * 8.0 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tflops_out = best;
    return SPARC_OK;
}

int sparc_layout_info(const SparcParams *p, int32_t hdr[4], int32_t *half_first, int32_t *half_cells,
                      int32_t *half_region, int max_halves)
{
    if (!p || !hdr || !half_first || !half_cells || !half_region) return fail(SPARC_E_INVALID, "null layout_info argument");
    if (p->n_segments < 2) return fail(SPARC_E_UNSUPPORTED, "n_segments < 2 is not supported");
    sparc::SparcLayout L;
    const Variant *v = pick_variant(p, &L);
    if (!v) return fail(SPARC_E_UNSUPPORTED, "n_segments too large (max 6657 wire segments)");
    const int nh = 64 * v->W;
    if (max_halves < nh) return fail(SPARC_E_INVALID, "half arrays too small (need 64 * warps per env)");
    hdr[0] = v->K; hdr[1] = v->W; hdr[2] = L.aligned; hdr[3] = L.hb[3];
    for (int h = 0; h < nh; h++) {
        int c0, n, r;
        sparc::half_geom(L, h, v->K / 2, c0, n, r);
        half_first[h] = c0; half_cells[h] = n; half_region[h] = r;
    }
    return SPARC_OK;
}

int sparc_launch_info(int n_envs, int n_segments, int32_t out[4])
{
    if (!out || n_envs <= 0 || n_segments < 2) return fail(SPARC_E_INVALID, "bad launch_info arguments");
    SparcParams q;                               // geometry only: no zone (the kernel variant does not depend on it
    memset(&q, 0, sizeof(q));                    // except through K when the aligned layout needs more slack)
    q.n_segments = n_segments;
    const Variant *v = pick_variant(&q, nullptr);
    if (!v) return fail(SPARC_E_UNSUPPORTED, "n_segments too large (max 6657 wire segments)");
    int blocks, threads;
    launch_geometry(v, n_envs, &blocks, &threads);
    out[0] = 32 * v->W;
    out[1] = v->K;
    out[2] = threads;
    out[3] = blocks;
    return SPARC_OK;
}

}  // extern "C"
