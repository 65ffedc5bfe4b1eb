// Kernel (b), latency mapping over thread-block clusters: ONE task per cluster of SIX CTAs (six SMs).
//
// Why: a handful of tasks (the reference's own sweep is 21 fits) leaves most of the GPU idle, and one task's run is a
// chain of ~4 000 dependent mini-batch steps.  Inside one SM more warps per task stop paying once the shared-memory
// pipe saturates (profiles/r2_*: 6 warps, 50 % shared pipe, 33 % issue).  Six SMs have six shared-memory pipes.
//
// Same arithmetic, same CANONICAL SUMMATION ORDER as fit_kernel.cu, so the weight trajectory is
// bit-identical to every other mapping and to the CPU oracle.  The canonical order splits the batch rows into six
// classes (position q mod 6, one fused chain per class and weight slot): CTA c of the cluster owns class c.
//   rows    CTA c walks its <= 8 rows of the mini-batch (q = c, c+6, ...), ONE row and one float2 column piece per
//           thread (96 threads = 8 rows x 12 pieces): zh -> h -> zy -> error -> dzy -> dzh, the same fused chains.
//   grads   CTA c forms ITS class sum p_c of every weight slot (4x4 tiles, one thread per tile, a chain over its rows).
//   reduce-scatter  the 48*LP weight slots are dealt to the six CTAs (8*LP each); every CTA sends the slots' class sums
//           to their owners' shared memory (st.async over DSMEM, completion counted on the owner's mbarrier).
//   update  the owner combines the six class sums in canonical order ((p0+p2+p4)+(p1+p3+p5)), applies Nadam (m, v in its
//           registers, two slots per thread), and
//   all-gather      writes the new weights into all six CTAs' weight copies (st.async again).
// Per step: two DSMEM exchanges (~4 KB in, ~4 KB out per CTA each), no cluster-wide barrier: an mbarrier per exchange,
// armed for its next phase BEFORE this CTA does anything that lets a peer start producing that phase.
// Validation is DEFERRED and runs beside training: the weights after an epoch's last update are the weights its validation
// needs AND the weights the next epoch's first step starts from, so three more warps per CTA (the validation warps) run the
// validation rows through the forward pass while the training warps do that first step's rows and gradients; each CTA's
// share of the epoch's squared errors (train, validation) rides along with the step's class sums (8 more bytes per peer),
// every CTA sums the six shares in rank order and takes the same early-stopping decision from the same bits BEFORE the
// step's update — a run that stops leaves the weights of its last epoch untouched.  Only the last epoch of a run that
// never stops needs a validation pass and an exchange of its own.
//
// Every wait is bounded (a clock64 deadline, then a trap): a protocol error would end the launch with an error, not hang the GPU.
// DSMEM here is this GPU's shared memory only; nothing in this file touches another GPU's memory.
#include "common.cuh"
#include "fit_params.cuh"

#include <cstdlib>

namespace hfrep {

constexpr int CL = 6;                 // CTAs per cluster = row classes of the canonical order
constexpr int RTH = 96;               // threads of a role: 8 rows x 12 float2 column pieces
constexpr int CTH = 2 * RTH;          // threads per CTA: three training warps + three validation warps
constexpr int RPC = 8;                // rows of a mini-batch pass per CTA (batch <= 48)

// ---- PTX helpers (cluster, mbarrier, asynchronous DSMEM stores)
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned cluster_rank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ unsigned map_to_rank(unsigned local_addr, unsigned rank) {
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arm(unsigned bar, unsigned tx_bytes) {      // one arrival + the bytes this phase expects
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(tx_bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// (default semantics: acquire at CTA scope.  The data this CTA waits for are asynchronous stores INTO ITS OWN shared memory,
// made visible by the barrier's transaction count — the same contract as a TMA load; a cluster-scope acquire would add an
// L1 invalidation (CCTL.IVALL) to every wait.)
__device__ __forceinline__ bool mbar_try(unsigned bar, unsigned parity) {
    unsigned ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void st_async_v4(unsigned raddr, float4 v, unsigned rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];"
                 ::"r"(raddr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(rbar) : "memory");
}
__device__ __forceinline__ void st_async_v2(unsigned raddr, float2 v, unsigned rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
                 ::"r"(raddr), "f"(v.x), "f"(v.y), "r"(rbar) : "memory");
}

template <int LP>
struct CLay {
    static constexpr int HS = (LP % 8 == 4) ? LP : LP + 4;
    static constexpr int CT = LP / 4;
    static constexpr int NT5 = 12 * CT;            // weight tiles (4x4): dWd (CT x 6) then dWe (6 x CT); <= 72 <= CTH
    static constexpr int NQ = 12 * LP;             // slot quads, tile-major [i][tile] as in fit_kernel.cu
    static constexpr int QPC = NQ / CL;            // quads owned by a CTA (2*LP)
    static_assert(NT5 <= RTH && 2 * QPC <= RTH, "one tile / one owned slot pair per training thread");
    __host__ __device__ static size_t floats(int T, int PS) {
        return (size_t)(T + 1) * XS                // Xs: the training block, row T all-zero
               + (size_t)EP * LP + (size_t)LP * WDS          // We, Wd
               + 3 * (size_t)(RPC + 1) * HS + (size_t)(RPC + 1) * XS   // H, dZh, H of the validation warps, dZy (row RPC = scratch)
               + (size_t)CL * QPC * 4              // part_in [source][owned quad][4]
               + NQ                                // woff
               + 2 * (size_t)((T + 3) & ~3)        // xoff tables
               + 2 * CL + 32 + 16                  // epoch mailbox, reduction scratch, barriers + flags
               + PS;
    }
};

template <int LP>
__global__ void __launch_bounds__(CTH, 1) fit_cluster_kernel(const FitParams p) {
    extern __shared__ __align__(16) float smem[];
    using Y = CLay<LP>;
    constexpr int HS = Y::HS, CT = Y::CT, NT5 = Y::NT5, NQ = Y::NQ, QPC = Y::QPC;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned c = cluster_rank();                       // this CTA's row class and slot share
    const int task = blockIdx.x / CL;
    const int T = p.T, E = p.E;
    const int L = p.task_L[task];

    // ---- carve-up
    float* Xs = smem;
    float* We = Xs + (size_t)(T + 1) * XS;
    float* Wd = We + EP * LP;
    float* H = Wd + LP * WDS;
    float* dZh = H + (RPC + 1) * HS;
    float* Hv = dZh + (RPC + 1) * HS;                        // the validation warps' H
    float* dZy = Hv + (RPC + 1) * HS;
    float* part_in = dZy + (RPC + 1) * XS;                   // [CL][QPC][4]
    int* woff = reinterpret_cast<int*>(part_in + CL * QPC * 4);
    const int TP = (T + 3) & ~3;
    int* xoff_tab = woff + NQ;
    float2* eloss = reinterpret_cast<float2*>(xoff_tab + 2 * TP);       // [CL] (train, validation) squared-error shares
    float* red = reinterpret_cast<float*>(eloss + CL);                    // [32]
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(red + 32);   // [3]: partials, weights, epoch
    unsigned* keyS = reinterpret_cast<unsigned*>(red + 32 + 16);
    int* permS = reinterpret_cast<int*>(keyS) + p.PS / 2;
    const int zero_row = T * XS;
    const unsigned barP = smem_u32(bars), barW = smem_u32(bars + 1), barE = smem_u32(bars + 2);
    const unsigned TX_SLOTS = (CL - 1) * QPC * 16u;           // bytes a CTA receives per exchange: five peers' share
    const unsigned TX_EPOCH = (CL - 1) * 8u;                  // ... plus, when an epoch's losses ride along, 8 bytes per peer

    // ---- stage the block, the kernels, the slot table
    const int xb = p.task_x ? p.task_x[task] : 0;
    const float* xg = p.X + (size_t)xb * T * E;
    for (int idx = tid; idx < (T + 1) * XS; idx += CTH) {
        const int r = idx / XS, cc = idx - r * XS;
        Xs[idx] = (r < T && cc < E) ? __ldg(xg + (size_t)r * E + cc) : 0.f;
    }
    for (int qd = tid; qd < NQ; qd += CTH) {                  // tile-major slot quad [i][tile] -> offset of its 4 weights from We
        const int i = qd / NT5, t = qd - i * NT5;
        int o;
        if (t < 6 * CT) { const int lt = t / 6, et = t - lt * 6; o = EP * LP + (lt * 4 + i) * WDS + et * 4; }
        else { const int u = t - 6 * CT; const int lt = u / 6, et = u - lt * 6; o = (et * 4 + i) * LP + lt * 4; }
        woff[qd] = o;
    }
    {
        const float* w0 = p.W0 + (size_t)task * p.w_stride;
        for (int idx = tid; idx < EP * LP; idx += CTH) {
            const int e = idx / LP, l = idx - e * LP;
            We[idx] = (e < E && l < L) ? w0[e * L + l] : 0.f;
        }
        for (int idx = tid; idx < LP * WDS; idx += CTH) {
            const int l = idx / WDS, e = idx - l * WDS;
            Wd[idx] = (l < L && e < E) ? w0[E * L + l * E + e] : 0.f;
        }
    }
    // A barrier's phase needs one arrival per warp of THIS CTA (made after the warp has left the wait on the previous
    // phase — so a phase can never complete, and the parity flip back, while a warp is still polling the one before) plus
    // the bytes of the five peers' asynchronous stores.  The tx-count is signed: a peer's bytes may land before the
    // arrival that announces them.
    auto arm = [&](unsigned bar, unsigned tx_bytes) {
        __syncwarp();
        if (lane == 0) { if (warp == 0) mbar_arm(bar, tx_bytes); else mbar_arrive(bar); }
    };
    if (tid == 0) {
        mbar_init(barP, CTH / 32); mbar_init(barW, CTH / 32); mbar_init(barE, CTH / 32);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    arm(barP, TX_SLOTS); arm(barW, TX_SLOTS); arm(barE, TX_EPOCH);      // phase 0 of each
    cluster_sync_all();                                      // every CTA's barriers exist before anyone sends

    // bounded wait: about a second, then the launch is ended with an error rather than left hanging
    auto wait = [&](unsigned bar, unsigned parity) {
        const long long t0 = clock64();
        while (!mbar_try(bar, parity)) {
            if (clock64() - t0 > 2000000000ll) __trap();
        }
    };

    const int* perm_t = p.perm ? p.perm + (size_t)task * p.perm_task_stride : nullptr;
    const float alpha = p.alpha;
    const float omb1 = 1.f - p.b1, omb2 = 1.f - p.b2;
    const NadamK nk{p.b1, omb1, p.b2, omb2, p.lr, p.eps};

    // roles: warps 0-2 train, warps 3-5 validate.  thread -> (local row rl, column piece pc): a warp holds 8 rows x 4 pieces
    const bool val_role = warp >= RTH / 32;
    const int rwarp = val_role ? warp - RTH / 32 : warp;
    const int rl = lane & 7, pc = 4 * rwarp + (lane >> 3);
    float* Hrole = val_role ? Hv : H;
    auto role_sync = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(val_role ? 2 : 1), "n"(RTH) : "memory"); };
    // owned slot pair of this thread (update phase, training threads): quad uq of this CTA's share, half uh
    const bool owner = tid < 2 * QPC;
    const int uq = tid >> 1, uh = tid & 1;
    const int my_qd = (int)c * QPC + uq;                     // global slot quad
    float2 m2 = make_float2(0.f, 0.f), v2 = make_float2(0.f, 0.f);   // Nadam moments of the owned pair
    // remote addresses of the six CTAs' copies of: the owned pair's weights, the barriers, this CTA's mailbox slots
    unsigned peer_w[CL], peer_barW[CL], peer_barE[CL], peer_eloss[CL];       // (compile-time indices only: registers)
#pragma unroll
    for (int r = 0; r < CL; ++r) {
        peer_barW[r] = map_to_rank(barW, r); peer_barE[r] = map_to_rank(barE, r);
        peer_w[r] = owner ? map_to_rank(smem_u32(We + woff[owner ? my_qd : 0] + 2 * uh), r) : 0u;
        peer_eloss[r] = map_to_rank(smem_u32(eloss + c), r);
    }
    const unsigned my_part_row = smem_u32(part_in + (size_t)c * QPC * 4);       // my source row of a part_in (same offset in every CTA)

    // ---- rows phase: one row, one float2 piece per thread.  Returns the thread's share of the squared error.
    auto rows_phase = [&](const int* xoff, int nb, bool train, float coef) -> float {
        const int q = (int)c + CL * rl;                      // position of this thread's row in the segment
        const bool valid = q < nb;
        const float* xr = Xs + (valid ? xoff[q] : zero_row);
        const int rs = valid ? rl : RPC;                     // scratch row for idle threads
        float2 zh = make_float2(0.f, 0.f);
        if (pc < LP / 2) {                                   // latent piece: zh = x . We[:, 2pc..2pc+1], fused chain over the ETFs
            const float* Wl = We + 2 * pc;
            float2 ze = make_float2(0.f, 0.f), zo = make_float2(0.f, 0.f);      // chains over the even / the odd ETFs
#pragma unroll
            for (int e4 = 0; e4 < EP / 4; ++e4) {
                const float4 xv = ld4(xr + 4 * e4);
                ze = __ffma2_rn(make_float2(xv.x, xv.x), ld2(Wl + (4 * e4 + 0) * LP), ze);
                zo = __ffma2_rn(make_float2(xv.y, xv.y), ld2(Wl + (4 * e4 + 1) * LP), zo);
                ze = __ffma2_rn(make_float2(xv.z, xv.z), ld2(Wl + (4 * e4 + 2) * LP), ze);
                zo = __ffma2_rn(make_float2(xv.w, xv.w), ld2(Wl + (4 * e4 + 3) * LP), zo);
            }
            zh = make_float2(__fadd_rn(ze.x, zo.x), __fadd_rn(ze.y, zo.y));
            st2(Hrole + rs * HS + 2 * pc, make_float2(lrelu(zh.x, alpha), lrelu(zh.y, alpha)));
        }
        role_sync();
        // ETF piece: zy = h . Wd[:, 2pc..2pc+1], fused chain over the latent units
        float2 zy = make_float2(0.f, 0.f);
        {
            const float* Wy = Wd + 2 * pc;
            const float* hr = Hrole + rs * HS;
            float2 ze = make_float2(0.f, 0.f), zo = make_float2(0.f, 0.f);      // chains over the even / the odd latent units
#pragma unroll
            for (int l4 = 0; l4 < CT; ++l4) {
                const float4 hv = ld4(hr + 4 * l4);
                ze = __ffma2_rn(make_float2(hv.x, hv.x), ld2(Wy + (4 * l4 + 0) * WDS), ze);
                zo = __ffma2_rn(make_float2(hv.y, hv.y), ld2(Wy + (4 * l4 + 1) * WDS), zo);
                ze = __ffma2_rn(make_float2(hv.z, hv.z), ld2(Wy + (4 * l4 + 2) * WDS), ze);
                zo = __ffma2_rn(make_float2(hv.w, hv.w), ld2(Wy + (4 * l4 + 3) * WDS), zo);
            }
            zy = make_float2(__fadd_rn(ze.x, zo.x), __fadd_rn(ze.y, zo.y));
        }
        const float2 xv = ld2(xr + 2 * pc);
        const float dx = lrelu(zy.x, alpha) - xv.x, dy = lrelu(zy.y, alpha) - xv.y;
        const float sse = valid ? fmaf(dy, dy, dx * dx) : 0.f;
        if (!train) { role_sync(); return sse; }             // (H may be overwritten by the next chunk after this)
        st2(dZy + rs * XS + 2 * pc, make_float2(lrelu_grad(zy.x, coef * dx, alpha), lrelu_grad(zy.y, coef * dy, alpha)));
        role_sync();
        if (pc < LP / 2) {                                   // dzh of the own latent columns: S_even + S_odd per column
            float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
            const float* dr = dZy + rs * XS;
            const float* w0r = Wd + (2 * pc) * WDS;
            const float* w1r = w0r + WDS;
#pragma unroll
            for (int e4 = 0; e4 < EP / 4; ++e4) {
                const float4 d = ld4(dr + 4 * e4), wa = ld4(w0r + 4 * e4), wb = ld4(w1r + 4 * e4);
                a0 = __ffma2_rn(make_float2(d.x, d.y), make_float2(wa.x, wa.y), a0);
                a0 = __ffma2_rn(make_float2(d.z, d.w), make_float2(wa.z, wa.w), a0);
                a1 = __ffma2_rn(make_float2(d.x, d.y), make_float2(wb.x, wb.y), a1);
                a1 = __ffma2_rn(make_float2(d.z, d.w), make_float2(wb.z, wb.w), a1);
            }
            st2(dZh + rs * HS + 2 * pc, make_float2(lrelu_grad(zh.x, __fadd_rn(a0.x, a0.y), alpha),
                                                   lrelu_grad(zh.y, __fadd_rn(a1.x, a1.y), alpha)));
        }
        return sse;                                          // (the caller's CTA-wide barrier makes dZh visible to both roles)
    };

    // ---- grads phase + reduce-scatter: this CTA's class sum of every weight slot, sent to the slots' owners
    auto grads_and_send = [&](const int* xoff, int nb) {
        const int nrows = nb > (int)c ? (nb - (int)c + CL - 1) / CL : 0;       // rows of class c in this batch
        if (tid < NT5) {
            const int t = tid;
            const bool dec = t < 6 * CT;
            int a0, b0, bstr;
            if (dec) { const int lt = t / 6, et = t - lt * 6; a0 = (int)(H - Xs) + lt * 4; b0 = (int)(dZy - Xs) + et * 4; bstr = XS; }
            else { const int u = t - 6 * CT; const int lt = u / 6, et = u - lt * 6; a0 = et * 4; b0 = (int)(dZh - Xs) + lt * 4; bstr = HS; }
            float2 acc[4][2];
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = make_float2(0.f, 0.f);
#pragma unroll
            for (int r = 0; r < RPC; ++r) {                  // ascending positions q = c + 6 r: the class chain.  Fixed trip count
                const bool live = r < nrows;                 // (all loads issue up front); rows past the class read the zero row
                const int ao = dec ? r * HS : xoff[live ? (int)c + CL * r : 0];
                const float4 a = ld4(Xs + (live ? a0 + ao : zero_row));
                const float4 b = ld4(Xs + (live ? b0 + r * bstr : zero_row));
                const float av[4] = {a.x, a.y, a.z, a.w};
                const float2 b01 = make_float2(b.x, b.y), b23 = make_float2(b.z, b.w);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float2 aa = make_float2(av[i], av[i]);
                    acc[i][0] = __ffma2_rn(aa, b01, acc[i][0]);
                    acc[i][1] = __ffma2_rn(aa, b23, acc[i][1]);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int qd = i * NT5 + t;
                const int own = qd / QPC, slot = qd - own * QPC;
                const float4 v = make_float4(acc[i][0].x, acc[i][0].y, acc[i][1].x, acc[i][1].y);
                if (own == (int)c) st4(part_in + ((size_t)c * QPC + slot) * 4, v);
                else st_async_v4(map_to_rank(my_part_row + slot * 16u, own), v, map_to_rank(barP, own));
            }
        }
    };

    // ---- update of the owned pair + all-gather of the new weights (two weight slots per training thread)
    auto quad_gradient2 = [&]() -> float2 {                  // canonical combination of the six class sums
        const float* pq = part_in + (size_t)uq * 4 + 2 * uh;
        auto at = [&](int s2) { return ld2(pq + (size_t)s2 * QPC * 4); };
        auto add2 = [](float2 x, float2 y) { return make_float2(__fadd_rn(x.x, y.x), __fadd_rn(x.y, y.y)); };
        const float2 even = add2(add2(at(0), at(2)), at(4));
        const float2 odd = add2(add2(at(1), at(3)), at(5));
        return add2(even, odd);
    };
    auto update_and_send = [&](int step, const float4& sa, const float4& sb) {
        if (!owner) return;
        const float2 g = quad_gradient2();
        const int o = woff[my_qd] + 2 * uh;
        if (p.grad_out != nullptr && step == p.dump_step) {  // test hook: the summed gradient, unpadded
            float* go = p.grad_out + (size_t)task * p.w_stride;
            const float gv[2] = {g.x, g.y};
            for (int k = 0; k < 2; ++k) {
                if (o >= EP * LP) { const int r = (o - EP * LP) / WDS, cc = (o - EP * LP) - r * WDS + k; if (r < L && cc < E) go[E * L + r * E + cc] = gv[k]; }
                else { const int r = o / LP, cc = o - r * LP + k; if (r < E && cc < L) go[r * L + cc] = gv[k]; }
            }
        }
        const float3 rc = make_float3(rcp_refined(sa.x), rcp_refined(sa.y), rcp_refined(sa.z));
        float2 wv = ld2(We + o);
        nadam_moments(g.x, m2.x, v2.x, nk); nadam_moments(g.y, m2.y, v2.y, nk);
        unsigned lo = 0xffffffffu;
        float dx = nadam_decrement<true>(g.x, m2.x, v2.x, sa, sb.x, rc, nk, lo);
        float dy = nadam_decrement<true>(g.y, m2.y, v2.y, sa, sb.x, rc, nk, lo);
        if (lo < kNadamSmallBits) {
            const float4 d4 = nadam_decrement_checked4(make_float4(g.x, g.y, 0.f, 0.f), make_float4(m2.x, m2.y, 0.f, 0.f),
                                                       make_float4(v2.x, v2.y, 1.f, 1.f), sa, sb.x, nk);
            dx = d4.x; dy = d4.y;
        }
        wv = make_float2(__fsub_rn(wv.x, dx), __fsub_rn(wv.y, dy));
#pragma unroll
        for (int r = 0; r < CL; ++r) {
            if (r == (int)c) st2(We + o, wv);
            else st_async_v2(peer_w[r], wv, peer_barW[r]);
        }
    };

    // ---- training run (same segment loop as fit_kernel.cu, resident mode)
    const int n_fit = p.n_fit, n_val = p.n_val, batch = p.batch;
    const int n_tb = (n_fit + batch - 1) / batch, n_vb = (n_val + batch - 1) / batch;
    int step = 0, wait_cnt = 0, epochs_run = 0, st = 0, cur = 0;
    unsigned phP = 0, phW = 0, phE = 0;                      // phase parities of the three barriers
    float best = INFINITY;
    const bool dev_shuffle = p.task_seed != nullptr && perm_t == nullptr;
    const unsigned stream_id = dev_shuffle ? p.task_seed[task] : 0u;
    auto make_perm = [&](int epoch) {                        // every CTA derives the same order (see fit_kernel.cu)
        for (int i = tid; i < n_fit; i += CTH) {
            unsigned h = stream_id * 0x9E3779B1u ^ (unsigned)(epoch + 1) * 0x85EBCA77u ^ (unsigned)(i + 1) * 0xC2B2AE3Du;
            h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
            keyS[i] = h;
        }
        __syncthreads();
        for (int i = tid; i < n_fit; i += CTH) {
            const unsigned k = keyS[i];
            int rank = 0;
            for (int j = 0; j < n_fit; ++j) { const unsigned kj = keyS[j]; rank += (kj < k) || (kj == k && j < i); }
            permS[rank] = i;
        }
        __syncthreads();
    };
    const bool ident = perm_t == nullptr && !dev_shuffle;
    auto fill_order = [&](int* tab, int epoch) {
        for (int i = tid; i < n_fit; i += CTH) {
            const int row = perm_t ? __ldg(perm_t + (size_t)epoch * n_fit + i) : (dev_shuffle ? permS[i] : i);
            tab[i] = row * XS;
        }
    };
    if (dev_shuffle) make_perm(0);
    for (int i = n_fit + tid; i < T; i += CTH) xoff_tab[i] = xoff_tab[TP + i] = i * XS;
    fill_order(xoff_tab, 0);
    __syncthreads();

#ifdef HFREP_CLUSTER_TIMING            // cycles per phase of CTA 0's thread 0, printed at the end (profiles/README.md)
    long long tk[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tl = 0;
#define TICK(i) do { const long long now_ = clock64(); if ((i) > 0) tk[i] += now_ - tl; tl = now_; } while (0)
#else
#define TICK(i) do { } while (0)
#endif
    // squared-error shares of one epoch: this role's warps reduce theirs (train: `tr`, validation: `va`) into red[0..5]
    auto reduce_shares = [&](float tr, float va) {
        const float v = warp_sum(val_role ? va : tr);
        if (lane == 0) red[warp] = v;
    };
    auto own_share = [&]() { return make_float2((red[0] + red[1]) + red[2], (red[3] + red[4]) + red[5]); };
    // the validation rows through the forward pass (validation warps; same weights as the training step beside them)
    auto validate = [&]() -> float {
        float vs = 0.f;
        for (int vb = 0; vb < n_vb; ++vb) {
            const int lo = vb * batch, nb = min(batch, n_val - lo);
            vs += rows_phase(xoff_tab + n_fit + lo, nb, false, 0.f);      // (the validation rows' offsets are the same in both tables)
        }
        return vs;
    };
    // every CTA's shares are in: the epoch's losses, its history entry, the early-stopping rule.  Returns true to stop.
    auto close_epoch = [&](int epoch, float2 mine) -> bool {
        float loss = 0.f, val = 0.f;
#pragma unroll
        for (int r = 0; r < CL; ++r) { const float2 e2 = r == (int)c ? mine : eloss[r]; loss += e2.x; val += e2.y; }
        loss = loss / (float)(E * n_fit);
        val = n_val > 0 ? val / (float)(E * n_val) : NAN;
        if (c == 0 && tid == 0) {
            p.loss_hist[(size_t)task * p.max_epochs + epoch] = loss;
            p.val_hist[(size_t)task * p.max_epochs + epoch] = val;
        }
        epochs_run = epoch + 1;
        if (!isfinite(loss)) st = 1;
        if (p.early_stop && n_val > 0) {                     // EarlyStopping.on_epoch_end (App. A.6), identical on all six CTAs
            ++wait_cnt;
            if (val < best) { best = val; wait_cnt = 0; }
            if (wait_cnt >= p.patience && epoch > 0) return true;
        }
        return false;
    };

    float lsum = 0.f;                                        // training threads: squared error of the epoch being trained
    bool stopped = false;
    for (int epoch = 0; epoch < p.max_epochs && !stopped; ++epoch) {
#pragma unroll 1
        for (int seg = 0; seg < n_tb; ++seg) {
            const int lo = seg * batch;
            const int nb = min(batch, n_fit - lo);
            const int* xoff = xoff_tab + cur * TP + lo;
            const bool last = seg + 1 == n_tb;
            const bool closing = seg == 0 && epoch > 0;      // this step also validates and closes epoch - 1
            if (last && !ident) {                            // next epoch's order into the other table
                const int ne = epoch + 1 < p.max_epochs ? epoch + 1 : epoch;
                if (dev_shuffle) make_perm(ne);
                fill_order(xoff_tab + (cur ^ 1) * TP, ne);
            }
            float4 sa, sb;
            asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(sa.x), "=f"(sa.y), "=f"(sa.z), "=f"(sa.w) : "l"(p.sched + 2 * step));
            asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(sb.x), "=f"(sb.y), "=f"(sb.z), "=f"(sb.w) : "l"(p.sched + 2 * step + 1));
            TICK(0);
            if (!val_role) {
                const float sse = rows_phase(xoff, nb, true, 2.f / (float)(nb * E));
                if (closing) { reduce_shares(lsum, 0.f); lsum = 0.f; }
                lsum += sse;
            } else if (closing) {
                reduce_shares(0.f, validate());
            }
            __syncthreads();                                 // H, dZy, dZh of the batch rows are complete
            TICK(1);
            grads_and_send(xoff, nb);                        // (training threads)
            __syncthreads();                                 // the own class sums are in part_in, the shares in red
            float2 mine = make_float2(0.f, 0.f);
            if (closing) {
                mine = own_share();
                if (tid == 0) {
#pragma unroll
                    for (int r = 0; r < CL; ++r)
                        if (r != (int)c) st_async_v2(peer_eloss[r], mine, map_to_rank(barP, r));
                }
            }
            TICK(2);
            wait(barP, phP); phP ^= 1;                       // the five peers' sums of the slots this CTA owns (+ their shares)
            arm(barP, TX_SLOTS + (last ? TX_EPOCH : 0u));    // the next step closes an epoch iff this one ends one
            TICK(3);
            if (closing && close_epoch(epoch - 1, mine)) { stopped = true; break; }      // before the update: weights stay
            update_and_send(step, sa, sb);
            __syncthreads();                                 // own weights written
            TICK(4);
            wait(barW, phW); phW ^= 1;                       // all six shares of the new weights are here
            arm(barW, TX_SLOTS);
            TICK(5);
            ++step;
            if (last && !ident) cur ^= 1;
            __syncthreads();
        }
    }
    if (!stopped) {                                          // the last epoch of a run that never stopped: its own validation pass
        TICK(0);
        reduce_shares(lsum, val_role ? validate() : 0.f);
        __syncthreads();
        const float2 mine = own_share();
        if (tid == 0) {
#pragma unroll
            for (int r = 0; r < CL; ++r)
                if (r != (int)c) st_async_v2(peer_eloss[r], mine, peer_barE[r]);
        }
        wait(barE, phE); phE ^= 1;
        close_epoch(p.max_epochs - 1, mine);
        TICK(6);
    }

#ifdef HFREP_CLUSTER_TIMING
    if (tid == 0 && blockIdx.x < CL)
        printf("cluster timing LP=%d cta=%u steps=%d epochs=%d: rows %lld grads+send %lld waitP %lld update+send %lld waitW %lld (cycles/step)\n",
               LP, c, step, epochs_run, tk[1] / max(step, 1), tk[2] / max(step, 1), tk[3] / max(step, 1), tk[4] / max(step, 1),
               tk[5] / max(step, 1));
#endif
    if (c == 0) {
        for (int e2 = epochs_run + tid; e2 < p.max_epochs; e2 += CTH) {
            p.loss_hist[(size_t)task * p.max_epochs + e2] = NAN;
            p.val_hist[(size_t)task * p.max_epochs + e2] = NAN;
        }
        float* wo = p.W_out + (size_t)task * p.w_stride;
        for (int idx = 2 * E * L + tid; idx < p.w_stride; idx += CTH) wo[idx] = 0.f;
        for (int idx = tid; idx < E * L; idx += CTH) { const int e = idx / L, l = idx - e * L; wo[idx] = We[e * LP + l]; }
        for (int idx = tid; idx < L * E; idx += CTH) { const int l = idx / E, e = idx - l * E; wo[E * L + idx] = Wd[l * WDS + e]; }
        if (tid == 0) {
            p.epochs_run[task] = epochs_run;
            p.status[task] = st;
        }
    }
    cluster_sync_all();                                      // no CTA leaves while a peer could still address its shared memory
}

// ---------------------------------------------------------------------------------- host side
template <int LP>
static int launch_cluster(const FitParams& p, cudaStream_t stream) {
    const size_t smem = sizeof(float) * CLay<LP>::floats(p.T, p.PS);
    HFREP_REQUIRE(smem <= 227 * 1024, "fit (cluster mapping): %zu bytes of shared memory needed", smem);
    auto kern = fit_cluster_kernel<LP>;
    HFREP_CUDA_OK(allow_max_dynamic_smem(kern));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)p.n_tasks * CL, 1, 1);
    cfg.blockDim = dim3(CTH, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    HFREP_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, p));
    return HFREP_OK;
}

// shapes the cluster kernel takes: a resident training block per task, mini-batches of at most 48 rows (8 per CTA)
bool fit_cluster_takes(const FitParams& p, int L_max) {
    const int LP = (L_max + 3) & ~3;
    if (LP < 4 || LP > 24 || p.batch > CL * RPC || p.force_tile) return false;
    const size_t fl = (size_t)(p.T + 1) * XS + 4096 + 2 * (size_t)((p.T + 3) & ~3) + p.PS;   // generous bound on CLay::floats
    return sizeof(float) * fl <= 200 * 1024;
}

int fit_cluster_dispatch(const FitParams& p, int L_max, cudaStream_t stream) {
    switch ((L_max + 3) & ~3) {
        case 4: return launch_cluster<4>(p, stream);
        case 8: return launch_cluster<8>(p, stream);
        case 12: return launch_cluster<12>(p, stream);
        case 16: return launch_cluster<16>(p, stream);
        case 20: return launch_cluster<20>(p, stream);
        case 24: return launch_cluster<24>(p, stream);
    }
    return set_err(HFREP_ERR_ARG, "fit (cluster mapping): L_max=%d outside 1..%d", L_max, HFREP_MAX_L);
}

}  // namespace hfrep
