// Kernel (b): batched autoencoder fit — one task (E -> L -> E bias-free LeakyReLU autoencoder,
// MSE, Nadam, validation tail, early stopping) per team of NW warps, whole training run in one launch.
//
// Replaces Autoencoder (Autoencoder_encapsulate.py:19-35) and AE.train (:72-96) of the reference,
// i.e. the Keras fit loop; semantics restated in SURVEY.md App. A.1-A.6.
//
// Per mini-batch step, three phases separated by team barriers:
//   rows    every warp of the team walks all batch rows, three rows per lane (lane pair g: rows g, g+16, g+32 of a
//           48-row pass), and owns float2 column pieces of the latent and the ETF block (2*NW lane classes).  The
//           rows' zh -> h -> zy -> y, squared error, dzy -> dzh accumulate in REGISTERS; one weight LDS.64 feeds
//           three rows (3 FFMA2).  h, dzy, dzh rows are parked in shared memory for the other classes and for the
//           weight gradients.
//   grads   dWd = H^T dZy and dWe = X_b^T dZh as 4x4 register tiles (outer products, 2 LDS.128 per 8 FFMA2), one
//           lane per tile; the tiles form groups of <= 32 and the batch rows are split over the warps of a group
//           (split-K); partial tiles go to shared memory in tile-major order.
//   update  over the 12*LP slot quads (four consecutive columns of a weight row): sum the partials in the canonical
//           order, Nadam (m, v in registers; shared memory for one-warp teams), write the weights back.
// CANONICAL SUMMATION ORDER (oracle/fitcore.py and oracle/c/fit_oracle.c follow the same one): every sum
// that feeds the weights has one fixed order in every mapping, so a run's weight trajectory is bit-identical across the
// mappings and to the CPU oracle — a LeakyReLU pre-activation within rounding of zero would otherwise send two correct
// float32 implementations down different branches (3e-5 jumps in the loss history).
//   zh[r][l] = Z_even + Z_odd, fused chains over the even / the odd ETFs in ascending order; zy[r][e] likewise over the
//   latent units   (padding adds exact zeros)
//   dzy = lrelu'(zy) * (coef * (y - x)), coef = 2 / (rows * E) rounded once
//   dh[r][l] = S_even + S_odd, fused chains over the even / the odd ETFs, ascending
//   dWd, dWe = (p0 + p2 + p4) + (p1 + p3 + p5), p_c = fused chain over the batch positions q = c, c+6, c+12, ...
//   Nadam in Keras' operation order, IEEE division and square root, every operation rounded on its own.
// This file is compiled with -fmad=false: the only fused operations are the explicit fma intrinsics.
// Shared memory (floats, rows 16-byte aligned).  Per CTA: Xs [(T+1)][28] (shared by the CTA's teams when all
// tasks train on the same block; cols >= E zero, row T zero), woff [12*LP] (tile-major slot quad -> weight offset).
// Per team: We [24][LP], Wd [LP][28], H [BP+1][HS], dZh [BP+1][HS], dZy [BP+1][28] (row BP = scratch row for
// idle lanes), part [KS][48*LP] (KS <= NW partial sums per slot, Lay::ks), mv [12*LP][2] float4 (one-warp teams),
// xoff [2][T] (position in the epoch's order -> row offset in Xs, this epoch's and the next one's: the fused
// window gather — shuffled orders cost nothing extra).
// LP = L rounded up to 4; zero padding is exactly neutral (zero activations, zero gradients, Nadam(0)=0).
#include "common.cuh"
#include "fit_params.cuh"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <type_traits>

namespace hfrep {

#ifndef HFREP_CHAIN_UNROLL
#define HFREP_CHAIN_UNROLL 4
#endif
constexpr int kChainUnroll = HFREP_CHAIN_UNROLL;   // a row class has 8 rows in a 48-row batch, 5 in the 30-row one
// (EP, XS, WDS, FitParams: fit_params.cuh)

template <int LP>
struct Lay {
    static constexpr int HS = (LP % 8 == 4) ? LP : LP + 4;
    static constexpr int CT = LP / 4;
    static constexpr int NT5 = 12 * CT;        // weight tiles: dWd (CT x 6) then dWe (6 x CT)
    static constexpr int NEL = NT5 * 16;       // = 48*LP weight slots, tile-major
    // grads phase: the tiles form TG groups of <= 32 (one lane per tile).  When the team's warps divide evenly over the
    // groups, a warp keeps ONE group and takes every KS-th batch row (KS = NW/TG partial sums per slot); otherwise every
    // warp walks all groups and takes every NW-th row (KS = NW).
    static constexpr int TG = (NT5 + 31) / 32;
    __host__ __device__ static constexpr bool grouped(int NW) { return TG > 1 && NW % TG == 0; }
    __host__ __device__ static constexpr int ks(int NW) { return grouped(NW) ? NW / TG : NW; }
    // The batch positions form six classes (q mod 6, one fused chain each: the canonical order); a warp takes the
    // classes kq, kq+KS, ...  KS is 1, 2, 3 or 6 for the team sizes 1, 2, 3, 6.  Partial sums kept per weight slot in
    // shared memory: the canonical combination (p0+p2+p4)+(p1+p3+p5) is formed by the warp itself when it owns a whole
    // parity (KS = 1: the total; KS = 2: one parity each), else all six class sums are stored.
    __host__ __device__ static constexpr int npart(int NW) { return ks(NW) >= 3 ? 6 : ks(NW); }
    // Nadam's m, v: registers when the team has >= 2 warps (at most 5 slot quads per lane), else shared memory
    __host__ __device__ static constexpr bool mv_in_regs(int NW) { return NW >= 2; }
    // XT (batch-tile mode): the training block stays in global memory / L2 and only the current mini-batch's rows
    // sit in a per-team tile [BP+1][XS], refilled with cp.async; else the whole block is resident once per CTA.
    // PRIV: resident mode with the training block per team instead of per CTA (tasks with their own blocks)
    __host__ __device__ static size_t team_floats(int BP, int NW, bool XT, int PS, int T, bool PRIV = false) {
        return (PRIV ? (size_t)(T + 1) * XS : 0) + (size_t)EP * LP + (size_t)LP * WDS + (size_t)(BP + 1) * HS * 2 + (size_t)(BP + 1) * XS +
               (size_t)npart(NW) * NEL + (mv_in_regs(NW) ? 0 : 2 * (size_t)NEL) + (XT ? 0 : 2 * (size_t)((T + 3) & ~3)) + 64 +
               (XT ? (size_t)(BP + 1) * XS : 0) + PS;
    }
    __host__ __device__ static size_t cta_floats(int T, bool XT, bool PRIV = false) { return ((XT || PRIV) ? 0 : (size_t)(T + 1) * XS) + NEL / 4; }
};

template <int NW>
__device__ __forceinline__ void team_sync(int team) {
    if constexpr (NW == 1) {
        __syncwarp();
    } else {
        asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "n"(NW * 32) : "memory");
    }
}

template <int NW>
__device__ __forceinline__ float team_sum(float v, float* red, int team, int tl) {
    v = warp_sum(v);
    if constexpr (NW == 1) {
        return v;
    } else {
        if ((tl & 31) == 0) red[tl >> 5] = v;
        team_sync<NW>(team);
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < NW; ++w) s += red[w];
        team_sync<NW>(team);
        return s;
    }
}



// LP: padded latent width; NW: warps per team (1, 2, 3 or 6: 2*NW lane classes must divide the 12 ETF float2 pieces);
// TPC: most teams per CTA (the launch picks blockDim.x / (32*NW) <= TPC teams from the shared-memory budget);
// FULL: fully unroll the k-loops of the rows phase (latency mapping: a handful of tasks, registers are free, the
// compiler software-pipelines the shared loads) instead of rolling them (throughput mapping);
// XT: batch-tile mode (see Lay::team_floats) — any block length T, any block per task, several teams per CTA
template <int LP, int NW, int TPC, bool FULL, bool XT>
__global__ void __launch_bounds__(32 * NW * TPC) fit_kernel(const FitParams p) {
    extern __shared__ __align__(16) float smem[];
    constexpr int NTH = 32 * NW;
    constexpr int HS = Lay<LP>::HS;
    constexpr int CT = Lay<LP>::CT;
    constexpr int NT5 = Lay<LP>::NT5;
    constexpr int NEL = Lay<LP>::NEL;
    const int team = threadIdx.x / NTH;
    const int tl = threadIdx.x - team * NTH;       // lane within the team
    const int warp = tl >> 5, lane = tl & 31;
    const int task = blockIdx.x * (blockDim.x / NTH) + team;
    const bool active = task < p.n_tasks;
    const int T = p.T, E = p.E, BP = p.BP;

    const bool priv = !XT && p.private_x;            // this team's own resident block
    const size_t cf = Lay<LP>::cta_floats(T, XT, priv);
    const size_t tf = Lay<LP>::team_floats(BP, NW, XT, p.PS, T, priv);
    // [Xs (resident mode only) | woff | team 0 | team 1 ...]
    int* woff = reinterpret_cast<int*>(smem + ((XT || priv) ? 0 : (size_t)(T + 1) * XS));
    float* base = smem + cf + (size_t)team * tf;
    float* We = base + (priv ? (size_t)(T + 1) * XS : 0);
    float* Wd = We + EP * LP;
    float* H = Wd + LP * WDS;
    float* dZh = H + (BP + 1) * HS;
    float* dZy = dZh + (BP + 1) * HS;
    float* part = dZy + (BP + 1) * XS;
    constexpr bool MVREG = Lay<LP>::mv_in_regs(NW);
    constexpr int NQ = NEL / 4;                    // slot quads: four consecutive columns of one weight row
    constexpr int QPL = MVREG ? (NQ + NTH - 1) / NTH : 1;   // quads per lane when m, v live in registers
    constexpr int KS = Lay<LP>::ks(NW);            // row-class stride of a warp in the grads phase (see Lay::npart)
    constexpr int NPART = Lay<LP>::npart(NW);      // partial sums per weight slot in shared memory
    static_assert(6 % KS == 0, "canonical order: the warps of a tile group must divide the six row classes");
    float4* mv = reinterpret_cast<float4*>(part + NPART * NEL);   // [NQ][2]: m, v of a quad (shared-memory variant)
    // resident mode: row-offset tables of a whole epoch (position in the epoch's order -> offset of the row in Xs: the
    // fused window gather, so shuffled orders cost nothing), [2][TP], one being read while the next epoch's is written
    const int TP = XT ? 0 : (T + 3) & ~3;
    int* xoff_tab = reinterpret_cast<int*>(part + NPART * NEL + (MVREG ? 0 : 2 * NEL));
    float* red = reinterpret_cast<float*>(xoff_tab + 2 * TP);          // [2][32], double-buffered per epoch
    // Xs: what the phases address rows in — the resident block (row r at xoff[r]) or the team's batch tile (row r at r*XS)
    float* Xs = XT ? red + 64 : (priv ? base : smem);
    unsigned* keyS = reinterpret_cast<unsigned*>(red + 64 + (XT ? (BP + 1) * XS : 0));   // [n_fit] shuffle keys
    int* permS = reinterpret_cast<int*>(keyS) + p.PS / 2;                                  // [n_fit] this epoch's row order
    const int zero_row = XT ? BP * XS : T * XS;      // all-zero row for idle lanes

    // ---- stage the training block (fused window load: rows of the scaled block -> padded tile)
    const int xb = (!active || !p.task_x) ? 0 : p.task_x[task];
    const float* xg = p.X + (size_t)xb * T * E;        // this task's training block in global memory
    {
        if constexpr (XT) {
            for (int idx = tl; idx < (BP + 1) * XS; idx += NTH) Xs[idx] = 0.f;   // pad columns and the zero row stay 0
        } else {
            const int n = (T + 1) * XS;
            const int i0 = priv ? tl : threadIdx.x, di = priv ? NTH : blockDim.x;     // own block: the team stages it
            for (int idx = i0; idx < n; idx += di) {
                const int r = idx / XS, c = idx - r * XS;
                Xs[idx] = (r < T && c < E) ? __ldg(xg + (size_t)r * E + c) : 0.f;
            }
        }
        // tile-major slot quad [i][tile] -> offset of its four consecutive weights from We
        for (int qd = threadIdx.x; qd < NQ; qd += blockDim.x) {
            const int i = qd / NT5, t = qd - i * NT5;
            int o;
            if (t < 6 * CT) {                            // dWd tile (lt, et): Wd[lt*4+i][et*4 ..]
                const int lt = t / 6, et = t - lt * 6;
                o = EP * LP + (lt * 4 + i) * WDS + et * 4;
            } else {                                     // dWe tile (lt, et): We[et*4+i][lt*4 ..]
                const int u = t - 6 * CT;
                const int lt = u / 6, et = u - lt * 6;
                o = (et * 4 + i) * LP + lt * 4;
            }
            woff[qd] = o;
        }
    }
    if (active) {
        const int L0 = p.task_L[task];
        const float* w0 = p.W0 + (size_t)task * p.w_stride;
        for (int idx = tl; idx < EP * LP; idx += NTH) {
            const int e = idx / LP, l = idx - e * LP;
            We[idx] = (e < E && l < L0) ? w0[e * L0 + l] : 0.f;
        }
        for (int idx = tl; idx < LP * WDS; idx += NTH) {
            const int l = idx / WDS, e = idx - l * WDS;
            Wd[idx] = (l < L0 && e < E) ? w0[E * L0 + l * E + e] : 0.f;
        }
        if constexpr (!MVREG)
            for (int idx = tl; idx < 2 * NQ; idx += NTH) mv[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    if (!active) return;
    const int L = p.task_L[task];

    const int* perm_t = p.perm ? p.perm + (size_t)task * p.perm_task_stride : nullptr;
    const float alpha = p.alpha;
    const float omb1 = 1.f - p.b1, omb2 = 1.f - p.b2;

    // ---------------------------------------------------------------- rows phase
    // Forward (and backward to dZh) of the batch rows.  Returns the lane's share of the squared error.  Lanes
    // without a row work on the all-zero row and park their (all-zero) results in scratch row BP, so the phase
    // needs no predication.  All multiply-adds are packed FFMA2 (fma.rn.f32x2, sm_100): two IEEE-rounded FP32
    // FMAs per issue slot, bit-identical to scalar fmaf.
    // Shared-memory loads, not FMAs, bound a one-row-per-lane layout: a broadcast LDS.128 costs 2 cycles of the SM's
    // shared-memory pipe and would feed only 2 FFMA2 (profiles/microbench/lds_wavefronts.cu: LDS.64 / LDS.128 run at
    // 1 / 2 cycles only when lane pairs share an address or the warp alternates between two).  So a lane holds
    // THREE rows (g, g+16, g+32 of the pass; g = lane/2) and float2 column pieces p = class + j*NC of the latent and
    // the ETF block (class = 2*warp + lane&1, NC = 2*NW): one LDS.64 (two-address pattern, 1 cycle) feeds 3 FFMA2,
    // row loads are pair-shared LDS.128 (2 cycles).  Every warp walks all rows, so h and dzy rows cross warps:
    // two team barriers per pass (a __syncwarp for one-warp teams).
    // Pieces past the block (LP/2 not a multiple of NC) are computed on whatever follows in shared memory and dropped.
    auto rows_pass = [&](auto RC, const int* xoff, int r0, int nb, bool train, float coef, float2& sse2) {
        constexpr int R = decltype(RC)::value;      // rows per lane in this pass: 3, or 2 when rows g+32 are all past the batch
        constexpr int NC = 2 * NW;
        constexpr int NPL = (LP / 2 + NC - 1) / NC;   // latent pieces per lane
        constexpr int NPE = (EP / 2) / NC;            // ETF pieces per lane (12 splits evenly over 2, 4, 6, 12 classes)
        static_assert((EP / 2) % NC == 0, "column-split mapping: 2*NW must divide 12");
        const int cls = 2 * warp + (lane & 1);
        const int g = lane >> 1;
        const float* Wl = We + 2 * cls;               // piece i of a We row at Wl + 2*NC*i
        const float* Wy = Wd + 2 * cls;               // piece i of a Wd row at Wy + 2*NC*i
        {
            const float* xr[R];
            int rs[R];
#pragma unroll
            for (int j = 0; j < R; ++j) {
                const int r = r0 + g + 16 * j;
                const bool valid = r < nb;
                xr[j] = Xs + (valid ? (XT ? r * XS : xoff[r]) : zero_row);
                rs[j] = valid ? r : BP;
            }
            // ---- zh pieces = x . We[:, own pieces]
            // two fused chains per output, over the even and over the odd ETFs, then their sum (canonical order)
            float2 zh[R][NPL];
            {
                float2 zc[2][R][NPL];
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int j = 0; j < R; ++j)
#pragma unroll
                        for (int i = 0; i < NPL; ++i) zc[h][j][i] = make_float2(0.f, 0.f);
#pragma unroll (FULL ? EP / 4 : 2)
                for (int e4 = 0; e4 < EP / 4; ++e4) {
                    float xe[R][4];
#pragma unroll
                    for (int j = 0; j < R; ++j) {
                        const float4 v = ld4(xr[j] + 4 * e4);
                        xe[j][0] = v.x; xe[j][1] = v.y; xe[j][2] = v.z; xe[j][3] = v.w;
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k)
#pragma unroll
                        for (int i = 0; i < NPL; ++i) {
                            const float2 w = ld2(Wl + (4 * e4 + k) * LP + 2 * NC * i);
#pragma unroll
                            for (int j = 0; j < R; ++j)
                                zc[k & 1][j][i] = __ffma2_rn(make_float2(xe[j][k], xe[j][k]), w, zc[k & 1][j][i]);
                        }
                }
#pragma unroll
                for (int j = 0; j < R; ++j)
#pragma unroll
                    for (int i = 0; i < NPL; ++i)
                        zh[j][i] = make_float2(__fadd_rn(zc[0][j][i].x, zc[1][j][i].x), __fadd_rn(zc[0][j][i].y, zc[1][j][i].y));
            }
#pragma unroll
            for (int i = 0; i < NPL; ++i)
                if (cls + NC * i < LP / 2) {
#pragma unroll
                    for (int j = 0; j < R; ++j)
                        st2(H + rs[j] * HS + 2 * (cls + NC * i), make_float2(lrelu(zh[j][i].x, alpha), lrelu(zh[j][i].y, alpha)));
                }
            team_sync<NW>(team);
            // ---- zy pieces = h . Wd[:, own pieces]
            float2 zy[R][NPE];
            {
                float2 zc[2][R][NPE];          // chains over the even / the odd latent units
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int j = 0; j < R; ++j)
#pragma unroll
                        for (int i = 0; i < NPE; ++i) zc[h][j][i] = make_float2(0.f, 0.f);
#pragma unroll (FULL ? CT : 2)
                for (int l4 = 0; l4 < CT; ++l4) {
                    float hl[R][4];
#pragma unroll
                    for (int j = 0; j < R; ++j) {
                        const float4 v = ld4(H + rs[j] * HS + 4 * l4);
                        hl[j][0] = v.x; hl[j][1] = v.y; hl[j][2] = v.z; hl[j][3] = v.w;
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k)
#pragma unroll
                        for (int i = 0; i < NPE; ++i) {
                            const float2 w = ld2(Wy + (4 * l4 + k) * WDS + 2 * NC * i);
#pragma unroll
                            for (int j = 0; j < R; ++j)
                                zc[k & 1][j][i] = __ffma2_rn(make_float2(hl[j][k], hl[j][k]), w, zc[k & 1][j][i]);
                        }
                }
#pragma unroll
                for (int j = 0; j < R; ++j)
#pragma unroll
                    for (int i = 0; i < NPE; ++i)
                        zy[j][i] = make_float2(__fadd_rn(zc[0][j][i].x, zc[1][j][i].x), __fadd_rn(zc[0][j][i].y, zc[1][j][i].y));
            }
            // ---- squared error and dLoss/dzy on the lane's pieces
#pragma unroll
            for (int j = 0; j < R; ++j)
#pragma unroll
                for (int i = 0; i < NPE; ++i) {
                    const float2 xv = ld2(xr[j] + 2 * (cls + NC * i));
                    const float dx = lrelu(zy[j][i].x, alpha) - xv.x, dy = lrelu(zy[j][i].y, alpha) - xv.y;
                    sse2.x = fmaf(dx, dx, sse2.x);
                    sse2.y = fmaf(dy, dy, sse2.y);
                    zy[j][i] = make_float2(lrelu_grad(zy[j][i].x, coef * dx, alpha), lrelu_grad(zy[j][i].y, coef * dy, alpha));
                }
            if (train) {
#pragma unroll
                for (int j = 0; j < R; ++j)
#pragma unroll
                    for (int i = 0; i < NPE; ++i) st2(dZy + rs[j] * XS + 2 * (cls + NC * i), zy[j][i]);
                team_sync<NW>(team);
                // ---- dzh pieces = (dzy . Wd[l, :]) * lrelu'(zh) for the own latent columns; the ETF index is the outer
                // loop (a float4 of each dzy row and of each own Wd row per step), two packed partial sums per (row, column)
                float2 acc[R][2 * NPL];
#pragma unroll
                for (int j = 0; j < R; ++j)
#pragma unroll
                    for (int c = 0; c < 2 * NPL; ++c) acc[j][c] = make_float2(0.f, 0.f);
#pragma unroll (FULL ? EP / 4 : 2)
                for (int e4 = 0; e4 < EP / 4; ++e4) {
                    float2 d01[R], d23[R];
#pragma unroll
                    for (int j = 0; j < R; ++j) {
                        const float4 v = ld4(dZy + rs[j] * XS + 4 * e4);
                        d01[j] = make_float2(v.x, v.y); d23[j] = make_float2(v.z, v.w);
                    }
#pragma unroll
                    for (int c = 0; c < 2 * NPL; ++c) {
                        const float4 w = ld4(Wd + (2 * (cls + NC * (c >> 1)) + (c & 1)) * WDS + 4 * e4);
#pragma unroll
                        for (int j = 0; j < R; ++j) {
                            acc[j][c] = __ffma2_rn(d01[j], make_float2(w.x, w.y), acc[j][c]);
                            acc[j][c] = __ffma2_rn(d23[j], make_float2(w.z, w.w), acc[j][c]);
                        }
                    }
                }
#pragma unroll
                for (int i = 0; i < NPL; ++i)
                    if (cls + NC * i < LP / 2) {
#pragma unroll
                        for (int j = 0; j < R; ++j)
                            st2(dZh + rs[j] * HS + 2 * (cls + NC * i),
                                make_float2(lrelu_grad(zh[j][i].x, acc[j][2 * i].x + acc[j][2 * i].y, alpha),
                                            lrelu_grad(zh[j][i].y, acc[j][2 * i + 1].x + acc[j][2 * i + 1].y, alpha)));
                    }
            }
        }
    };
    auto rows_phase = [&](const int* xoff, int nb, bool train, float coef) -> float {
        float2 sse2 = make_float2(0.f, 0.f);
#pragma unroll 1
        for (int r0 = 0; r0 < nb; r0 += 48) {
            if (nb - r0 <= 32) rows_pass(std::integral_constant<int, 2>(), xoff, r0, nb, train, coef, sse2);
            else rows_pass(std::integral_constant<int, 3>(), xoff, r0, nb, train, coef, sse2);
        }
        return sse2.x + sse2.y;
    };

    // ---------------------------------------------------------------- weight-gradient phase (row classes over warps)
    auto grads_phase = [&](const int* xoff, int nb) {
        constexpr bool GRP = Lay<LP>::grouped(NW);
        constexpr int TPG = NT5 / Lay<LP>::TG;          // tiles per group (12*CT/TG: a multiple of 6)
        const int grp = GRP ? warp % Lay<LP>::TG : 0;
        const int kq = GRP ? warp / Lay<LP>::TG : warp;  // first row class of this warp
        const int tend = GRP ? (grp + 1) * TPG : NT5;
#pragma unroll 1
        for (int t = grp * TPG + lane; t < tend; t += 32) {
            const bool dec = t < 6 * CT;
            int a0, b0, bstr;
            if (dec) {                                   // dWd tile: A = H[:, lt*4..], B = dZy[:, et*4..]
                const int lt = t / 6, et = t - lt * 6;
                a0 = (int)(H - Xs) + lt * 4; b0 = (int)(dZy - Xs) + et * 4; bstr = XS;
            } else {                                     // dWe tile: A = X_b[:, et*4..], B = dZh[:, lt*4..]
                const int u = t - 6 * CT;
                const int lt = u / 6, et = u - lt * 6;     // lt-major like dWd: six consecutive lanes share a chunk
                a0 = et * 4; b0 = (int)(dZh - Xs) + lt * 4; bstr = HS;
            }
            // the fused chain of one row class: positions c, c+6, c+12, ...
            auto class_chain = [&](int c, float2 (&acc)[4][2]) {
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = make_float2(0.f, 0.f);
#pragma unroll kChainUnroll
                for (int q = c; q < nb; q += 6) {
                    const int ao = dec ? q * HS : (XT ? q * XS : xoff[q]);
                    const float4 a = ld4(Xs + a0 + ao);
                    const float4 b = ld4(Xs + b0 + q * bstr);
                    const float av[4] = {a.x, a.y, a.z, a.w};
                    const float2 b01 = make_float2(b.x, b.y), b23 = make_float2(b.z, b.w);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float2 aa = make_float2(av[i], av[i]);
                        acc[i][0] = __ffma2_rn(aa, b01, acc[i][0]);
                        acc[i][1] = __ffma2_rn(aa, b23, acc[i][1]);
                    }
                }
            };
            auto add_to = [&](float2 (&tot)[4][2], const float2 (&acc)[4][2]) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int h = 0; h < 2; ++h)
                        tot[i][h] = make_float2(__fadd_rn(tot[i][h].x, acc[i][h].x), __fadd_rn(tot[i][h].y, acc[i][h].y));
            };
            auto store = [&](int slot, const float2 (&v)[4][2]) {
                float* dst = part + slot * NEL;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    st4(dst + i * (NT5 * 4) + t * 4, make_float4(v[i][0].x, v[i][0].y, v[i][1].x, v[i][1].y));   // lanes -> consecutive 16 B
            };
            // One copy of the chain code per kernel: the classes of this warp are walked by run-time loops.
            float2 acc[4][2];
            if constexpr (KS >= 3) {                     // every class sum goes to shared memory; the update phase combines
#pragma unroll 1
                for (int c = kq; c < 6; c += KS) {
                    class_chain(c, acc);
                    store(c, acc);
                }
            } else {                                     // a parity's sum (p_c + p_c+2) + p_c+4.  KS = 2: this warp's parity.
                float2 par[4][2];                        // KS = 1: both, the even sum parked in the slot meanwhile
#pragma unroll 1
                for (int p0 = kq; p0 < 2; p0 += KS) {
#pragma unroll 1
                    for (int c = p0; c < 6; c += 2) {
                        class_chain(c, acc);
                        if (c == p0) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) { par[i][0] = acc[i][0]; par[i][1] = acc[i][1]; }
                        } else {
                            add_to(par, acc);
                        }
                    }
                    if (p0 != kq) {                      // T_even + T_odd
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float4 ev = ld4(part + kq * NEL + i * (NT5 * 4) + t * 4);
                            acc[i][0] = make_float2(ev.x, ev.y); acc[i][1] = make_float2(ev.z, ev.w);
                        }
                        add_to(par, acc);
                    }
                    store(kq, par);
                }
            }
        }
    };

    // ---------------------------------------------------------------- update phase (Nadam, element-wise)
    // A lane handles slot quads (four consecutive columns of one weight row): the NW partial sums, the weights and
    // (one-warp teams) m, v move as 128-bit shared accesses; the partials are summed in warp order.
    float4 m4r[QPL], v4r[QPL];
#pragma unroll
    for (int j = 0; j < QPL; ++j) m4r[j] = v4r[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    auto add4 = [](const float4& a, const float4& b) {
        return make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), __fadd_rn(a.w, b.w));
    };
    auto quad_gradient = [&](int qd) -> float4 {        // canonical combination of the class sums (see Lay::npart)
        const float* pq = part + 4 * qd;
        if constexpr (NPART == 6) {
            const float4 even = add4(add4(ld4(pq), ld4(pq + 2 * NEL)), ld4(pq + 4 * NEL));
            const float4 odd = add4(add4(ld4(pq + NEL), ld4(pq + 3 * NEL)), ld4(pq + 5 * NEL));
            return add4(even, odd);
        } else if constexpr (NPART == 2) {
            return add4(ld4(pq), ld4(pq + NEL));
        } else {
            return ld4(pq);
        }
    };
    // Nadam in Keras' operation order with IEEE-exact division and square root (common.cuh: nadam_delta).  The fast
    // sequences run unconditionally; if the smallest non-zero operand of the quad is below 2^-64 (never seen: gradients
    // and moments of these fits stay above 1e-14) the quad is redone with the compiler's range-checked operations.
    const NadamK nk{p.b1, omb1, p.b2, omb2, p.lr, p.eps};
    auto update_quad = [&](int qd, float4& m4, float4& v4, const float4& sa, const float4& sb, const float3& rc) {
        const float4 g = quad_gradient(qd);
        const int o = woff[qd];
        float4 wv = ld4(We + o);
        nadam_moments(g.x, m4.x, v4.x, nk); nadam_moments(g.y, m4.y, v4.y, nk);
        nadam_moments(g.z, m4.z, v4.z, nk); nadam_moments(g.w, m4.w, v4.w, nk);
        unsigned lo = 0xffffffffu;
        float4 d;
        d.x = nadam_decrement<true>(g.x, m4.x, v4.x, sa, sb.x, rc, nk, lo);
        d.y = nadam_decrement<true>(g.y, m4.y, v4.y, sa, sb.x, rc, nk, lo);
        d.z = nadam_decrement<true>(g.z, m4.z, v4.z, sa, sb.x, rc, nk, lo);
        d.w = nadam_decrement<true>(g.w, m4.w, v4.w, sa, sb.x, rc, nk, lo);
        if (lo < kNadamSmallBits) d = nadam_decrement_checked4(g, m4, v4, sa, sb.x, nk);
        wv.x = __fsub_rn(wv.x, d.x); wv.y = __fsub_rn(wv.y, d.y); wv.z = __fsub_rn(wv.z, d.z); wv.w = __fsub_rn(wv.w, d.w);
        st4(We + o, wv);
    };
    // test hook (cfg.dump_step): the summed gradient of one step, unpadded, to global memory.  Kept out of the update
    // loop: a branch per quad there would stop the compiler from interleaving the quads' dependency chains.
    auto dump_gradient = [&]() {
        float* go = p.grad_out + (size_t)task * p.w_stride;
        for (int qd = tl; qd < NQ; qd += NTH) {
            const float4 g = quad_gradient(qd);
            const float gv[4] = {g.x, g.y, g.z, g.w};
            const int o = woff[qd];
            for (int k = 0; k < 4; ++k) {
                if (o >= EP * LP) {
                    const int r = (o - EP * LP) / WDS, c = (o - EP * LP) - r * WDS + k;
                    if (r < L && c < E) go[E * L + r * E + c] = gv[k];
                } else {
                    const int r = o / LP, c = o - r * LP + k;
                    if (r < E && c < L) go[r * L + c] = gv[k];
                }
            }
        }
    };
    auto update_phase = [&](int step, const float4& sa, const float4& sb) {
        if (p.grad_out != nullptr && step == p.dump_step) dump_gradient();
        const float3 rc = make_float3(rcp_refined(sa.x), rcp_refined(sa.y), rcp_refined(sa.z));
        if constexpr (MVREG) {
#pragma unroll
            for (int j = 0; j < QPL; ++j)
                if (NQ % NTH == 0 || tl + j * NTH < NQ) update_quad(tl + j * NTH, m4r[j], v4r[j], sa, sb, rc);
        } else {
#pragma unroll 2
            for (int qd = tl; qd < NQ; qd += NTH) {
                float4 m4 = mv[2 * qd], v4 = mv[2 * qd + 1];
                update_quad(qd, m4, v4, sa, sb, rc);
                mv[2 * qd] = m4; mv[2 * qd + 1] = v4;
            }
        }
    };

    // ---------------------------------------------------------------- training run
    // An epoch is a list of segments: the train mini-batches, then the validation chunks.  One loop (one
    // call site per phase keeps the code small and the phases inlined).  The row-offset table of segment
    // s+1 is written while segment s runs (double buffer), so a train step costs three team barriers and
    // a validation chunk one.
    const int n_fit = p.n_fit, n_val = p.n_val, batch = p.batch;
    const int n_tb = (n_fit + batch - 1) / batch, n_vb = (n_val + batch - 1) / batch, n_seg = n_tb + n_vb;
    int step = 0, wait = 0, epochs_run = 0, st = 0, cur = 0;
    float best = INFINITY;
    // Device-side stand-in for Keras' shuffle=True: the row order of an epoch is the stable argsort of 32-bit hash keys
    // of (task stream, epoch, row) — a counter-based generator, so no state and no host table; oracle: fitcore.device_perm.
    const bool dev_shuffle = p.task_seed != nullptr && perm_t == nullptr;
    const unsigned stream_id = dev_shuffle ? p.task_seed[task] : 0u;
    auto make_perm = [&](int epoch) {
        for (int i = tl; i < n_fit; i += NTH) {
            unsigned h = stream_id * 0x9E3779B1u ^ (unsigned)(epoch + 1) * 0x85EBCA77u ^ (unsigned)(i + 1) * 0xC2B2AE3Du;
            h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
            keyS[i] = h;
        }
        team_sync<NW>(team);
        for (int i = tl; i < n_fit; i += NTH) {
            const unsigned k = keyS[i];
            int rank = 0;
            for (int j = 0; j < n_fit; ++j) {
                const unsigned kj = keyS[j];
                rank += (kj < k) || (kj == k && j < i);
            }
            permS[rank] = i;
        }
        team_sync<NW>(team);
    };
    // batch-tile mode: refill the tile with the rows of a segment
    auto set_rows = [&](int epoch, int seg) {
        const bool val = seg >= n_tb;
        const int lo = (val ? seg - n_tb : seg) * batch;
        const int nb = min(batch, (val ? n_val : n_fit) - lo);
        // cp.async of the rows' E floats (16-, 8- or 4-byte pieces by alignment)
        const int pw = ((E & 3) == 0 && (((size_t)xb * T * E) & 3) == 0) ? 4 : (((E & 1) == 0 && (((size_t)xb * T * E) & 1) == 0) ? 2 : 1);
        const int ppr = E / pw;                       // pieces per row
        for (int idx = tl; idx < nb * ppr; idx += NTH) {
            const int q = idx / ppr, c = (idx - q * ppr) * pw;
            const int row = val ? n_fit + lo + q : (perm_t ? __ldg(perm_t + (size_t)epoch * n_fit + lo + q) : (dev_shuffle ? permS[lo + q] : lo + q));
            const float* src = xg + (size_t)row * E + c;
            const unsigned dst = (unsigned)__cvta_generic_to_shared(Xs + q * XS + c);
            if (pw == 4) asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
            else if (pw == 2) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
            else asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // resident mode: the training part of an epoch's offset table (the validation tail never changes)
    const bool ident = perm_t == nullptr && !dev_shuffle;   // the same order every epoch: one table, written once
    auto fill_order = [&](int* tab, int epoch) {
        for (int i = tl; i < n_fit; i += NTH) {
            const int row = perm_t ? __ldg(perm_t + (size_t)epoch * n_fit + i) : (dev_shuffle ? permS[i] : i);
            tab[i] = row * XS;
        }
    };
    auto rows_ready = [&]() {                          // the tile refill issued by this thread has landed
        if constexpr (XT) asm volatile("cp.async.wait_group 0;" ::: "memory");
    };
    if (dev_shuffle) make_perm(0);
    if constexpr (XT) {
        set_rows(0, 0);
    } else {
        for (int i = n_fit + tl; i < T; i += NTH) xoff_tab[i] = xoff_tab[TP + i] = i * XS;
        fill_order(xoff_tab, 0);
    }
    rows_ready();
    team_sync<NW>(team);
    for (int epoch = 0; epoch < p.max_epochs; ++epoch) {
        float lsum = 0.f, vsum = 0.f;
#pragma unroll 1
        for (int seg = 0; seg < n_seg; ++seg) {
            const bool train = seg < n_tb;
            const int lo = (train ? seg : seg - n_tb) * batch;
            const int nb = min(batch, (train ? n_fit : n_val) - lo);
            const int* xoff = xoff_tab + cur * TP + (train ? lo : n_fit + lo);   // (unused in batch-tile mode)
            const bool last = seg + 1 == n_seg;
            const int ne = epoch + 1 < p.max_epochs ? epoch + 1 : epoch;
            // what the next segment needs.  Resident mode: nothing within an epoch; in its last segment the next epoch's
            // order goes to the other table.  Tile mode refills the single tile once its last reader is done.
            auto prefetch_next = [&]() {
                if constexpr (XT) {
                    if (!last) set_rows(epoch, seg + 1);
                    else {
                        if (dev_shuffle) make_perm(ne);   // this epoch's order is no longer needed: its last tile is out
                        set_rows(ne, 0);
                    }
                } else if (last && !ident) {
                    if (dev_shuffle) make_perm(ne);
                    fill_order(xoff_tab + (cur ^ 1) * TP, ne);
                }
            };
            if constexpr (!XT) prefetch_next();
            // this step's Nadam scalars: the global load is issued here and first used after two phases, so its latency
            // (the only global access of a step) is off the critical path of a one-team CTA
            float4 sa = make_float4(1.f, 1.f, 1.f, 0.f), sb = make_float4(0.f, 0.f, 0.f, 0.f);
            if (train) {
                asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];"
                             : "=f"(sa.x), "=f"(sa.y), "=f"(sa.z), "=f"(sa.w) : "l"(p.sched + 2 * step));
                asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];"
                             : "=f"(sb.x), "=f"(sb.y), "=f"(sb.z), "=f"(sb.w) : "l"(p.sched + 2 * step + 1));
            }
            float sse = 0.f;
            if (!(p.debug_skip & 1)) {
                sse = rows_phase(xoff, nb, train, 2.f / (float)(nb * E));
            }
            if (train) {
                lsum += sse;
                team_sync<NW>(team);
                if (!(p.debug_skip & 2)) grads_phase(xoff, nb);
                team_sync<NW>(team);
                if constexpr (XT) prefetch_next();          // the copy overlaps the update phase
                if (!(p.debug_skip & 4)) update_phase(step, sa, sb);
                ++step;
            } else {
                vsum += sse;
                if constexpr (XT) { team_sync<NW>(team); prefetch_next(); }
            }
            rows_ready();
            if (!XT && last && !ident) cur ^= 1;
            team_sync<NW>(team);
        }
        // epoch losses: one barrier; the scratch alternates between epochs so no second barrier is needed
        float loss, val;
        {
            const float l1 = warp_sum(lsum), v1 = warp_sum(vsum);
            if constexpr (NW == 1) {
                loss = l1; val = v1;
            } else {
                float* rd = red + (epoch & 1) * 32;
                if (lane == 0) { rd[warp] = l1; rd[8 + warp] = v1; }
                team_sync<NW>(team);
                loss = 0.f; val = 0.f;
#pragma unroll
                for (int w = 0; w < NW; ++w) { loss += rd[w]; val += rd[8 + w]; }
            }
        }
        loss = loss / (float)(E * n_fit);
        val = n_val > 0 ? val / (float)(E * n_val) : NAN;
        if (tl == 0) {
            p.loss_hist[(size_t)task * p.max_epochs + epoch] = loss;
            p.val_hist[(size_t)task * p.max_epochs + epoch] = val;
        }
        epochs_run = epoch + 1;
        if (!isfinite(loss)) st = 1;
        if (p.early_stop && n_val > 0) {         // EarlyStopping.on_epoch_end (App. A.6); with no validation rows there is
            ++wait;                              // no val_loss: Keras warns, leaves `wait` alone and trains every epoch
            if (val < best) { best = val; wait = 0; }
            if (wait >= p.patience && epoch > 0) break;
        }
    }

    // the caller's buffers need no initialisation: history entries past the last epoch run read NaN, the kernel row's tail
    // (the task's own L packs 2*E*L of the w_stride floats) reads zero
    for (int e2 = epochs_run + tl; e2 < p.max_epochs; e2 += NTH) {
        p.loss_hist[(size_t)task * p.max_epochs + e2] = NAN;
        p.val_hist[(size_t)task * p.max_epochs + e2] = NAN;
    }
    float* wo = p.W_out + (size_t)task * p.w_stride;
    for (int idx = 2 * E * L + tl; idx < p.w_stride; idx += NTH) wo[idx] = 0.f;
    for (int idx = tl; idx < E * L; idx += NTH) {
        const int e = idx / L, l = idx - e * L;
        wo[idx] = We[e * LP + l];
    }
    for (int idx = tl; idx < L * E; idx += NTH) {
        const int l = idx / E, e = idx - l * E;
        wo[E * L + idx] = Wd[l * WDS + e];
    }
    if (tl == 0) {
        p.epochs_run[task] = epochs_run;
        p.status[task] = st;
    }
}

// ---------------------------------------------------------------------------------- host side
constexpr size_t kMaxSmem = 227 * 1024;

template <int LP, int NW>
static size_t fit_smem_bytes(const FitParams& p, int tpc, bool xt) {
    const bool priv = !xt && p.private_x;
    return sizeof(float) * (Lay<LP>::cta_floats(p.T, xt, priv) + tpc * Lay<LP>::team_floats(p.BP, NW, xt, p.PS, p.T, priv));
}

// Resident mode: the training block sits in shared memory for the whole run — one copy shared by the CTA's teams when all
// tasks train on the same block (n_x == 1), one copy per team when every task has its own (private_x).  Batch-tile mode (a
// block too long for shared memory) stages only the current mini-batch's rows.  The launch
// takes the largest team count (<= TPC) whose shared memory fits.
template <int LP, int NW, int TPC, bool FULL, bool XT>
static int launch_fit_mode(const FitParams& p, cudaStream_t stream) {
    int tpc = TPC;
    if (const char* cap = getenv("HFREP_TPC")) tpc = std::max(1, std::min(TPC, atoi(cap)));   // tuning only
    while (tpc > 1 && fit_smem_bytes<LP, NW>(p, tpc, XT) > kMaxSmem) --tpc;
    // two half-size CTAs per SM beat one big one when both fit (finer scheduling granularity)
    if (tpc > 1 && tpc % 2 == 0 && 2 * fit_smem_bytes<LP, NW>(p, tpc / 2, XT) <= kMaxSmem) tpc /= 2;
    const size_t smem = fit_smem_bytes<LP, NW>(p, tpc, XT);
    HFREP_REQUIRE(smem <= kMaxSmem, "fit: %zu bytes of shared memory needed (batch=%d too large for the narrow kernel)", smem,
                  p.batch);
    auto kern = fit_kernel<LP, NW, TPC, FULL, XT>;
    HFREP_CUDA_OK(allow_max_dynamic_smem(kern));
    const int grid = (p.n_tasks + tpc - 1) / tpc;
    kern<<<grid, 32 * NW * tpc, smem, stream>>>(p);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

template <int LP, int NW, int TPC, bool FULL>
static int launch_fit(const FitParams& p, cudaStream_t stream) {
    FitParams q = p;
    q.private_x = !p.shared_x;                       // a block per task: resident per team when it fits, else batch tiles
    const bool resident = fit_smem_bytes<LP, NW>(q, 1, false) <= kMaxSmem && !p.force_tile &&
                          (p.shared_x || !getenv("HFREP_NO_PRIVATE_BLOCKS"));
    if (!resident) q.private_x = 0;
    return resident ? launch_fit_mode<LP, NW, TPC, FULL, false>(q, stream)
                    : launch_fit_mode<LP, NW, TPC, FULL, true>(q, stream);
}

// warps per task (every mapping runs the same phases; they differ in how many warps share a task's columns):
//   6  up to ~5 tasks per SM (one team per CTA, k-loops fully unrolled): the latency mapping;
//   2  the throughput mapping beyond that;
//   1, 3  kept for experiments: 1 has the least shared-memory traffic per task but needs more resident tasks than
//      shared memory admits to hide its latency; 3 sits between 2 and 6 and wins nowhere on the measured sweep.
template <int LP>
static int dispatch_mapping(const FitParams& p, int wpt, cudaStream_t stream) {
#ifndef HFREP_SLIM                                   // (HFREP_SLIM: experiment builds with the two main mappings only)
    if (wpt == 1) return launch_fit<LP, 1, 8, false>(p, stream);
    if (wpt == 3) return launch_fit<LP, 3, (LP <= 8 ? 8 : 6), false>(p, stream);
#endif
    if (wpt == 2) return launch_fit<LP, 2, 6, false>(p, stream);
    return launch_fit<LP, 6, 1, true>(p, stream);
}

int fit_dispatch(const FitParams& p, int L_max, int wpt, cudaStream_t stream) {
    const int LP = (L_max + 3) & ~3;
    switch (LP) {
#ifndef HFREP_SLIM
        case 4: return dispatch_mapping<4>(p, wpt, stream);
        case 12: return dispatch_mapping<12>(p, wpt, stream);
        case 16: return dispatch_mapping<16>(p, wpt, stream);
#endif
        case 8: return dispatch_mapping<8>(p, wpt, stream);
        case 20: return dispatch_mapping<20>(p, wpt, stream);
        case 24: return dispatch_mapping<24>(p, wpt, stream);
    }
    return set_err(HFREP_ERR_ARG, "fit: L_max=%d outside 1..%d", L_max, HFREP_MAX_L);
}

}  // namespace hfrep

using namespace hfrep;

extern "C" void hfrep_fit_cfg_default(hfrep_fit_cfg* c) {
    // literals of Autoencoder_encapsulate.py:19-35, :79-96 and Keras 2.7 Nadam() defaults
    memset(c, 0, sizeof(*c));
    c->E = 22; c->L_max = 8; c->T = 168; c->n_fit = 126; c->batch = 48; c->max_epochs = 1000;
    c->patience = 5; c->early_stop = 1; c->alpha = .2f;
    c->lr = 1e-3f; c->beta_1 = .9f; c->beta_2 = .999f; c->epsilon = 1e-7f; c->schedule_decay = .004f; c->decay_base = .96f;
    c->n_x = 1; c->warps_per_task = 0; c->dump_step = 0;
}

extern "C" int hfrep_nadam_schedule(const hfrep_fit_cfg* c, int n_steps, float* out) {
    HFREP_REQUIRE(c && out && n_steps >= 0, "nadam_schedule: null argument");
    // Keras 2.7 Nadam._prepare_local in float32
    float m_cache = 1.f;
    for (int s = 0; s < n_steps; ++s) {
        const float t = (float)(s + 1), nxt = (float)(s + 2);
        // float32 pow taken as the correctly rounded value (double pow, then rounded)
        auto pw = [](float x, float y) { return (float)pow((double)x, (double)y); };
        const float mu_t = c->beta_1 * (1.f - .5f * pw(c->decay_base, c->schedule_decay * t));
        const float mu_t1 = c->beta_1 * (1.f - .5f * pw(c->decay_base, c->schedule_decay * nxt));
        const float ms_new = m_cache * mu_t;
        const float ms_next = ms_new * mu_t1;
        float* o = out + (size_t)HFREP_SCHED_STRIDE * s;
        o[0] = 1.f - ms_new;                 // divides the gradient
        o[1] = 1.f - ms_next;                // divides the first moment
        o[2] = 1.f - pw(c->beta_2, t);       // divides the second moment
        o[3] = 1.f - mu_t;
        o[4] = mu_t1;
        o[5] = mu_t;
        o[6] = o[7] = 0.f;
        m_cache = ms_new;
    }
    return HFREP_OK;
}

static int fit_batched_impl(const hfrep_fit_cfg* c, const float* X, int n_tasks, const int32_t* task_L,
                            const int32_t* task_x, const float* W0, int w_stride, const int32_t* perm,
                            int64_t perm_task_stride, const uint32_t* task_seed, const float* sched, float* W_out,
                            float* loss_hist, float* val_hist, int32_t* epochs_run, int32_t* status, float* grad_out,
                            void* stream) {
    HFREP_REQUIRE(c && X && task_L && W0 && sched && W_out && loss_hist && val_hist && epochs_run && status,
                  "fit_batched: null argument");
    HFREP_REQUIRE(n_tasks >= 0, "fit_batched: n_tasks=%d", n_tasks);
    if (n_tasks == 0) return HFREP_OK;
    HFREP_REQUIRE(c->E >= 1 && c->E <= HFREP_MAX_E, "fit_batched: E=%d outside 1..%d", c->E, HFREP_MAX_E);
    HFREP_REQUIRE(c->L_max >= 1 && c->L_max <= HFREP_MAX_L, "fit_batched: L_max=%d outside 1..%d", c->L_max, HFREP_MAX_L);
    HFREP_REQUIRE(c->T >= 2 && c->n_fit >= 1 && c->n_fit <= c->T, "fit_batched: T=%d n_fit=%d", c->T, c->n_fit);
    HFREP_REQUIRE(c->batch >= 1 && c->batch <= 256, "fit_batched: batch=%d outside 1..256", c->batch);
    HFREP_REQUIRE(c->max_epochs >= 1, "fit_batched: max_epochs=%d", c->max_epochs);
    HFREP_REQUIRE(w_stride >= 2 * c->E * c->L_max, "fit_batched: w_stride=%d < 2*E*L_max", w_stride);
    HFREP_REQUIRE(c->n_x >= 1 && (c->n_x == 1 || task_x), "fit_batched: n_x=%d needs task_x", c->n_x);
    HFREP_REQUIRE(((uintptr_t)sched & 15) == 0, "fit_batched: sched must be 16-byte aligned");
    HFREP_REQUIRE(c->patience >= 0, "fit_batched: patience=%d", c->patience);

    FitParams p;
    p.X = X; p.task_L = task_L; p.task_x = task_x; p.W0 = W0; p.perm = perm; p.task_seed = task_seed;
    p.PS = (task_seed && !perm) ? ((2 * c->n_fit + 3) & ~3) : 0;
    p.sched = reinterpret_cast<const float4*>(sched);
    p.W_out = W_out; p.loss_hist = loss_hist; p.val_hist = val_hist; p.epochs_run = epochs_run; p.status = status;
    p.grad_out = grad_out; p.perm_task_stride = perm_task_stride;
    p.n_tasks = n_tasks; p.E = c->E; p.T = c->T; p.n_fit = c->n_fit; p.n_val = c->T - c->n_fit; p.batch = c->batch;
    p.BP = (c->batch + 3) & ~3; p.max_epochs = c->max_epochs; p.patience = c->patience; p.early_stop = c->early_stop;
    p.w_stride = w_stride; p.shared_x = (c->n_x == 1); p.dump_step = c->dump_step;
    { const char* dbg = getenv("HFREP_DEBUG_SKIP"); p.debug_skip = dbg ? atoi(dbg) : 0; }
    p.force_tile = getenv("HFREP_FORCE_TILE") != nullptr;
    p.private_x = 0;
    p.alpha = c->alpha; p.lr = c->lr; p.b1 = c->beta_1; p.b2 = c->beta_2; p.eps = c->epsilon;

    int wpt = c->warps_per_task;
    if (wpt == 0) {                                  // pick the mapping from the amount of parallel work in THIS call
        const int sms = sm_count();                  // (callers that split a job over several calls pass it explicitly)
        // measured on the L = 1..21 sweep x seeds (profiles/README.md): a six-CTA cluster per task while every cluster has
        // its own SMs; 6 warps win up to ~5 tasks per SM, 2 warps beyond (narrow tasks in large numbers: one warp per task
        // — L = 8 x 4096 tasks 6.6 vs 7.8 ms per 100 epochs; L = 20: 14.8 vs 12.1; with a training block per task (batch-tile
        // staging) 2 warps stay ahead: 100 000 path tasks 207 vs 219 ms)
        wpt = n_tasks * 6 <= sms ? HFREP_WPT_CLUSTER
                                 : (n_tasks <= 5 * sms ? 6 : ((c->L_max <= 8 && n_tasks >= 10 * sms && p.shared_x) ? 1 : 2));
    }
    HFREP_REQUIRE(wpt == 1 || wpt == 2 || wpt == 3 || wpt == 6 || wpt == HFREP_WPT_CLUSTER,
                  "fit_batched: warps_per_task=%d (0, 1, 2, 3, 6 or HFREP_WPT_CLUSTER)", wpt);
    if (wpt == HFREP_WPT_CLUSTER) {
        if (fit_cluster_takes(p, c->L_max) && !getenv("HFREP_NO_CLUSTER"))
            return fit_cluster_dispatch(p, c->L_max, (cudaStream_t)stream);
        wpt = 6;                                     // long blocks, batches over 48 rows: the one-CTA latency mapping
    }
    return fit_dispatch(p, c->L_max, wpt, (cudaStream_t)stream);
}

extern "C" int hfrep_fit_batched(const hfrep_fit_cfg* c, const float* X, int n_tasks, const int32_t* task_L,
                                 const int32_t* task_x, const float* W0, int w_stride, const int32_t* perm,
                                 int64_t perm_task_stride, const uint32_t* task_seed, const float* sched, float* W_out,
                                 float* loss_hist, float* val_hist, int32_t* epochs_run, int32_t* status, float* grad_out,
                                 void* stream) {
    return fit_batched_impl(c, X, n_tasks, task_L, task_x, W0, w_stride, perm, perm_task_stride, task_seed, sched, W_out,
                            loss_hist, val_hist, epochs_run, status, grad_out, stream);
}

// Rolling refit (Mode W): kernel (a) builds every window's MinMax-scaled training block into the caller's workspace,
// kernel (b) fits the tasks on them.
extern "C" size_t hfrep_fit_windows_workspace(int n_windows, int T_win, int E) {
    if (n_windows <= 0 || T_win <= 0 || E <= 0) return 0;
    return sizeof(float) * (size_t)n_windows * T_win * E;
}

extern "C" int hfrep_fit_windows(const hfrep_fit_cfg* c, const float* R, int T_all, int C, int col_lo,
                                 const int32_t* win_starts, int n_windows, int n_tasks, const int32_t* task_L,
                                 const int32_t* task_win, const float* W0, int w_stride, const int32_t* perm,
                                 int64_t perm_task_stride, const uint32_t* task_seed, const float* sched, void* ws,
                                 size_t ws_bytes, float* W_out, float* loss_hist, float* val_hist, int32_t* epochs_run,
                                 int32_t* status, void* stream) {
    HFREP_REQUIRE(c && R && win_starts && task_win, "fit_windows: null argument");
    HFREP_REQUIRE(n_windows >= 1, "fit_windows: n_windows=%d", n_windows);
    const size_t need = hfrep_fit_windows_workspace(n_windows, c->T, c->E);
    HFREP_REQUIRE(ws && ws_bytes >= need, "fit_windows: workspace of %zu bytes needed, %zu given", need, ws_bytes);
    HFREP_REQUIRE(((uintptr_t)ws & 15) == 0, "fit_windows: workspace must be 16-byte aligned");
    float* blocks = static_cast<float*>(ws);
    int rc = launch_window_gather(R, T_all, C, win_starts, n_windows, c->T, col_lo, c->E, /*minmax=*/1, blocks,
                                  (cudaStream_t)stream);
    if (rc != HFREP_OK) return rc;
    hfrep_fit_cfg cc = *c;
    cc.n_x = n_windows > 1 ? n_windows : 2;          // every task trains on its own block: one team per CTA
    return fit_batched_impl(&cc, blocks, n_tasks, task_L, task_win, W0, w_stride, perm, perm_task_stride, task_seed, sched,
                            W_out, loss_hist, val_hist, epochs_run, status, nullptr, stream);
}

// ------------------------------------------------------------------------------------------------------------
// hfrep_fit_windows_static: the rolling refit for callers that must not allocate device memory.
//
// Who that is: a host process that gave the whole GPU to its own arena at start-up (PyTorch's caching allocator
// with a memory fraction, an RMM pool, a risk engine with a fixed device-memory plan) has nothing left for a
// cudaMalloc made on its behalf, and a scheduler that keeps other streams busy cannot afford the device-wide
// synchronisation an allocation or a free implies.  Such a caller still owns its inputs and outputs, but it has
// nowhere to put the scaled window blocks that kernel (a) hands to kernel (b).  Here that scratch is a fixed
// array inside the library image: every GPU that loads libhfrep.so reserves HFREP_STATIC_SCRATCH_BYTES for it
// until the process exits.
//
// The array's name is only a handle on the host; its device address is looked up with cudaGetSymbolAddress the
// first time the entry point runs on a device (with that device current) and remembered per device.  Make that
// first call outside any stream capture — and before the host's arena claims what is left of the GPU: with CUDA's
// default lazy module loading the array (and each kernel variant's code) takes its device memory at that first use,
// not when the library is loaded (CUDA_MODULE_LOADING=EAGER moves it to load time).
// All calls on one device share the one array: two calls whose kernels
// overlap on the GPU (different streams, no dependency) overwrite each other's blocks.  Nothing in the array
// has to start at zero — kernel (a) rewrites every element the fit reads — so no clearing is done between calls.
// No cudaMalloc / cudaFree / allocating library routine is called on this path.
static __device__ __align__(256) float g_window_scratch[HFREP_STATIC_SCRATCH_BYTES / sizeof(float)];

static int static_scratch_address(float** out) {
    static std::atomic<float*> cached[64];       // per device; host threads may race to fill a slot with the same value
    int dev = -1;
    HFREP_CUDA_OK(cudaGetDevice(&dev));
    HFREP_REQUIRE(dev >= 0 && dev < 64, "fit_windows_static: device ordinal %d not supported", dev);
    float* ptr = cached[dev].load(std::memory_order_acquire);
    if (ptr == nullptr) {
        HFREP_CUDA_OK(cudaGetSymbolAddress((void**)&ptr, g_window_scratch));
        HFREP_REQUIRE(ptr != nullptr, "fit_windows_static: no device address for the scratch array");
        cached[dev].store(ptr, std::memory_order_release);
    }
    *out = ptr;
    return HFREP_OK;
}

extern "C" int hfrep_fit_windows_static(const hfrep_fit_cfg* c, const float* R, int T_all, int C, int col_lo,
                                        const int32_t* win_starts, int n_windows, int n_tasks,
                                        const int32_t* task_L, const int32_t* task_win, const float* W0, int w_stride,
                                        const int32_t* perm, int64_t perm_task_stride, const uint32_t* task_seed,
                                        const float* sched, float* W_out, float* loss_hist, float* val_hist,
                                        int32_t* epochs_run, int32_t* status, void* stream) {
    HFREP_REQUIRE(c != nullptr, "fit_windows_static: null argument");
    const size_t need = hfrep_fit_windows_workspace(n_windows, c->T, c->E);
    HFREP_REQUIRE(n_windows >= 1 && need > 0, "fit_windows_static: n_windows=%d T=%d E=%d", n_windows, c->T, c->E);
    HFREP_REQUIRE(need <= (size_t)HFREP_STATIC_SCRATCH_BYTES,
                  "fit_windows_static: %d windows x %d rows x %d columns need %zu bytes of scratch, the built-in array "
                  "holds %zu; use hfrep_fit_windows with your own workspace", n_windows, c->T, c->E, need,
                  (size_t)HFREP_STATIC_SCRATCH_BYTES);
    float* scratch = nullptr;
    const int rc = static_scratch_address(&scratch);
    if (rc != HFREP_OK) return rc;
    return hfrep_fit_windows(c, R, T_all, C, col_lo, win_starts, n_windows, n_tasks, task_L, task_win, W0, w_stride, perm,
                             perm_task_stride, task_seed, sched, scratch, (size_t)HFREP_STATIC_SCRATCH_BYTES, W_out, loss_hist,
                             val_hist, epochs_run, status, stream);
}
