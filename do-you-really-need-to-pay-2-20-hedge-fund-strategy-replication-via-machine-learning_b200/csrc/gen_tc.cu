// Tensor-core (tcgen05) path of the generator's recurrent contractions.
//   hfrep_tc_selftest: D[64 x 400] = A[64 x K] * B[400 x K]^T in 3xTF32 on one CTA — operands written by threads in the
//   canonical un-swizzled K-major layout, accumulators in tensor memory, read back with tcgen05.ld.  It pins the
//   descriptor fields, the M = 64 accumulator layout (row r -> lane 32*(r/16) + r%16) and the error of the 3xTF32 split
//   that the generator kernel builds on (tests/test_gpu_parity.py).
#include "common.cuh"
#include "tc.cuh"

#include <cstdlib>
#include <type_traits>

namespace hfrep {

constexpr int TCM = 64;            // rows (paths) per CTA
constexpr int TCN = 400;           // gate columns
constexpr int TCN0 = 256, TCN1 = TCN - TCN0;

__global__ void __launch_bounds__(128) tc_selftest_kernel(const float* A, const float* B, int K, float* D, int half) {
    extern __shared__ __align__(128) unsigned char smraw[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned char* a_hi = smraw;
    unsigned char* a_lo = a_hi + TCM * K * 4;
    unsigned char* b_hi = a_lo + TCM * K * 4;
    unsigned char* b_lo = b_hi + TCN * K * 4;

    if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
    if (tid == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
    for (int i = tid; i < TCM * K; i += blockDim.x) {
        const int r = i / K, k = i - r * K;
        float hi, lo;
        tc::split_tf32(A[i], hi, lo);
        *reinterpret_cast<float*>(a_hi + tc::kmajor_off(r, k, K)) = hi;
        *reinterpret_cast<float*>(a_lo + tc::kmajor_off(r, k, K)) = lo;
    }
    for (int i = tid; i < TCN * K; i += blockDim.x) {
        const int r = i / K, k = i - r * K;
        float hi, lo;
        tc::split_tf32(B[i], hi, lo);
        *reinterpret_cast<float*>(b_hi + tc::kmajor_off(r, k, K)) = hi;
        *reinterpret_cast<float*>(b_lo + tc::kmajor_off(r, k, K)) = lo;
    }
    tc::fence_async_smem();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    // half = 1: the accumulator in lanes 16-31 of every TMEM quarter (the other half-subpartition an M = 64 tile may take)
    const uint32_t tmem = tmem_slot + ((uint32_t)(16 * half) << 16);

    if (tid == 0) {
        const uint32_t sbo = (uint32_t)(K >> 2) * 128u, lbo = 128u;
        const uint32_t i0 = tc::idesc_tf32(TCM, TCN0), i1 = tc::idesc_tf32(TCM, TCN1);
        bool acc = false;
        for (int ks = 0; ks < K / 8; ++ks) {
            const uint32_t koff = (uint32_t)ks * 256u;
            const unsigned char* as[3] = {a_hi, a_lo, a_hi};
            const unsigned char* bs[3] = {b_hi, b_hi, b_lo};
            for (int t = 0; t < 3; ++t) {
                const uint64_t ad = tc::smem_desc(tc::smem_u32(as[t]) + koff, lbo, sbo);
                const uint64_t bd0 = tc::smem_desc(tc::smem_u32(bs[t]) + koff, lbo, sbo);
                const uint64_t bd1 = tc::smem_desc(tc::smem_u32(bs[t]) + koff + (TCN0 / 8) * sbo, lbo, sbo);
                tc::mma_tf32(tmem, ad, bd0, i0, acc);
                tc::mma_tf32(tmem + TCN0, ad, bd1, i1, acc);
                acc = true;
            }
        }
        tc::mma_commit(&bar);
    }
    tc::mbar_wait(&bar, 0);
    tc::tc_fence_after();
    // warp w reads TMEM lanes 32w .. 32w+31; rows 16w + lane live in its lanes 0..15
    for (int c0 = 0; c0 < TCN; c0 += 8) {
        float v[8];
        tc::tmem_ld8(tmem_slot + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0, v);
        if ((lane >> 4) == half) {
#pragma unroll
            for (int i = 0; i < 8; ++i) D[(size_t)(16 * warp + (lane & 15)) * TCN + c0 + i] = v[i];
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) tc::tmem_free<512>(tmem_slot);
}

// MMA issue-rate probe: every CTA's thread 0 issues `iters` kind::tf32 MMAs (M = 64, N as given, K = 8) on zeroed operands,
// commits, waits.  Timed from outside; tells what a small-N MMA costs next to a wide one (wide fit).
__global__ void __launch_bounds__(128) tc_mma_rate_kernel(int N, int iters) {
    extern __shared__ __align__(128) unsigned char smraw[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tc::tmem_alloc<256>(&tmem_slot);
    if (tid == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
    for (int i = tid; i < (64 + 256) * 8; i += blockDim.x) reinterpret_cast<float*>(smraw)[i] = 0.f;
    tc::fence_async_smem();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = tmem_slot;
    if (tid == 0) {
        const uint32_t idesc = tc::idesc_tf32(64, N);
        const uint64_t ad = tc::smem_desc(tc::smem_u32(smraw), 128u, 256u);
        const uint64_t bd = tc::smem_desc(tc::smem_u32(smraw) + 64 * 8 * 4, 128u, 256u);
        for (int i = 0; i < iters; ++i) tc::mma_tf32(tmem, ad, bd, idesc, i > 0);
        tc::mma_commit(&bar);
    }
    tc::mbar_wait(&bar, 0);
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) tc::tmem_free<256>(tmem);
}

// =====================================================================================================================
// The generator forward on tensor cores (units = 100: what GAN/MTSS_WGAN_GP.py:221-235 builds).
// One CTA carries 64 paths through all T steps.  Per step and LSTM layer the gate pre-activations
//   Z[64 x 400] = [x_t | h_{t-1}] [W; U]           (K = 40 + 104 for layer 1, 104 + 104 for layer 2, zero padded)
// are one chain of kind::tf32 MMAs (M = 64, N = 256 + 144, K = 8 each) in 3xTF32 — a_hi b_hi + a_lo b_hi + a_hi b_lo,
// FP32 accumulators in tensor memory — with
//   * the A operands (x_t / a1_t, h1, h2: TF32 hi and lo planes, canonical un-swizzled K-major layout) resident in shared
//     memory and rewritten in place by the epilogue threads every step,
//   * the B operands (the weights, split into hi / lo planes and laid out per K = 8 step by hfrep_generator_tc_prepare,
//     gate columns permuted so that a unit's four gates are adjacent) streamed from L2 through a two-stage ring of 25.6 KB
//     bulk copies (cp.async.bulk, THIS GPU's memory, no multicast), one mbarrier pair per stage,
//   * warps 0-7: epilogue (TMEM -> registers, gates, cell state in registers, LayerNorm, operands of the next MMA chain,
//     Dense, output), warp 8: MMA issue (one thread), warp 9: weight loader (one thread).
// M = 64 accumulators occupy lanes 0-15 of each 32-lane TMEM quarter; the epilogue warps hand half of what they load to
// their lanes 16-31 by shuffle, so every lane works on (row, 25 units): lane (row, q) owns units 4i + q, i < 25 (their
// operand columns are then compile-time offsets), and gate column n = 200*(q/2) + 8i + 4*(q%2) + g holds gate g of unit
// 4i + q, so that one 32-column TMEM load brings four units of both lane halves.
// =====================================================================================================================
constexpr int GU = 100, GG = 400, GKX = 40, GKH = 104;
constexpr int G_NKS1 = (GKX + GKH) / 8, G_NKS2 = (GKH + GKH) / 8, G_NKS = G_NKS1 + G_NKS2;       // 18, 26, 44 K-steps per time step
constexpr uint32_t G_PLANE_B = GG * 8 * 4, G_STAGE = 2 * G_PLANE_B;                               // 12 800, 25 600 bytes
constexpr uint32_t G_SBO_B = 256, G_LBO = 128;
constexpr uint32_t G_SBO_A = (GKH / 4) * 128, G_PLANE_A = 8 * G_SBO_A;                            // 3 328, 26 624 bytes
constexpr int G_ND = 40, G_NKD = GKH / 8;                                                       // Dense: N = 40 (36 + pad), 13 K-steps over a2
constexpr uint32_t G_PLANE_D = G_ND * 8 * 4, G_BLOCK_D = 2 * G_PLANE_D;                          // 1 280, 2 560 bytes
constexpr int G_EPI = 256;                                                                       // epilogue threads
constexpr int G_UPL = 25;                                                                        // units per lane

struct GenTcParams {
    const float* W;        // packed weights (hfrep.h): biases, LayerNorm, Dense read from here
    const unsigned char* tcw;   // prepared B blocks: [G_NKS][hi plane | lo plane] of the LSTM kernels, then [G_NKD] of the Dense kernel
    const float* noise; float* out;
    int n, T, F_in, F_out, lrelu1, lrelu2;
    float alpha, ln_eps;
};

// 2^x; results below the normal range flush to zero (1 + e is 1 either way)
__device__ __forceinline__ float ex2_fast(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float rcp_fast(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// gate column n -> (unit, gate): n = 200*pr + 8*i + 4*hh + g  <->  unit 4*i + 2*pr + hh
__host__ __device__ constexpr int gate_col_unit(int n) { return 4 * ((n % 200) >> 3) + 2 * (n / 200) + ((n >> 2) & 1); }
__host__ __device__ constexpr int gate_col_gate(int n) { return n & 3; }
__device__ __forceinline__ uint32_t a_off(int row, int k) {            // element (row, k) of a 64 x 104 operand plane
    return (uint32_t)(row >> 3) * G_SBO_A + (uint32_t)(k >> 2) * 128u + (uint32_t)(row & 7) * 16u + (uint32_t)(k & 3) * 4u;
}
__device__ __forceinline__ void epi_sync() { asm volatile("bar.sync 1, %0;" ::"n"(G_EPI) : "memory"); }

__global__ void __launch_bounds__(320, 1) generator_tc_kernel(const GenTcParams p) {
    extern __shared__ __align__(128) unsigned char smraw[];
    __shared__ __align__(8) uint64_t bar_full[2], bar_empty[2], bar_z, bar_a;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned char* H1 = smraw;                          // [hi | lo] planes of h1
    unsigned char* AX = H1 + 2 * G_PLANE_A;             // x_t during layer 1's MMAs, a1_t during layer 2's
    unsigned char* H2 = AX + 2 * G_PLANE_A;
    unsigned char* ring = H2 + 2 * G_PLANE_A;           // 2 stages of B
    float* b1p = reinterpret_cast<float*>(ring + 2 * G_STAGE);    // [400] permuted gate biases
    float* b2p = b1p + GG;
    float* lng = b2p + GG;                              // gamma1, beta1, gamma2, beta2 [100 each]
    float* lnA = lng + 4 * GU;                          // [64][4]
    float* lnB = lnA + 256;                             // [64][4]

    const int U = GU, G = GG, F_in = p.F_in, F_out = p.F_out, T = p.T;
    const float* W1 = p.W;
    const float* b1 = W1 + (size_t)F_in * G + (size_t)U * G;
    const float* g1 = b1 + G;
    const float* W2 = g1 + 2 * U;
    const float* b2 = W2 + 2 * (size_t)U * G;
    const float* g2 = b2 + G;
    const float* Kd = g2 + 2 * U;
    const float* bdg = Kd + (size_t)U * F_out;

    if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
    if (tid == 0) {
        tc::mbar_init(&bar_full[0], 1); tc::mbar_init(&bar_full[1], 1);
        tc::mbar_init(&bar_empty[0], 1); tc::mbar_init(&bar_empty[1], 1);
        tc::mbar_init(&bar_z, 1); tc::mbar_init(&bar_a, G_EPI / 32);
        tc::fence_barrier_init();
    }
    for (int i = tid; i < G; i += blockDim.x) {
        const int u = gate_col_unit(i), g = gate_col_gate(i);
        b1p[i] = __ldg(b1 + g * U + u);
        b2p[i] = __ldg(b2 + g * U + u);
    }
    for (int i = tid; i < 4 * U; i += blockDim.x) {               // gamma1, beta1, gamma2, beta2, each stored [q][i] for unit 4i + q
        const int arr = i / U, r = i - arr * U, qq = r / G_UPL, ii = r - qq * G_UPL;
        lng[i] = __ldg((arr < 2 ? g1 : g2) + (arr & 1) * U + 4 * ii + qq);
    }
    for (int i = tid; i < (int)(6 * G_PLANE_A / 4); i += blockDim.x) reinterpret_cast<float*>(smraw)[i] = 0.f;   // h1 = h2 = 0, pads
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = tmem_slot;
    const int path0 = blockIdx.x * 64;

    if (warp < G_EPI / 32) {
        // ============================================================ epilogue threads
        const int s = warp & 3, pr = warp >> 2, l16 = lane & 15, hh = lane >> 4;
        const int row = 16 * s + l16, q = 2 * pr + hh;             // this lane: units 4i + q, i < 25
        const int path = path0 + row;
        const bool live = path < p.n;
        const uint32_t trow = tmem + ((uint32_t)(32 * s) << 16);
        const float alpha = p.alpha;
        float c1[G_UPL], c2[G_UPL];
#pragma unroll
        for (int j = 0; j < G_UPL; ++j) c1[j] = c2[j] = 0.f;
        unsigned par_z = 0;
        const int kx0 = 10 * q;                                        // this thread's 10 columns of the x operand
        auto put = [&](unsigned char* region, int k, float v) {
            float hi, lo;
            tc::split_tf32_trunc(v, hi, lo);
            const uint32_t o = a_off(row, k);
            *reinterpret_cast<float*>(region + o) = hi;
            *reinterpret_cast<float*>(region + G_PLANE_A + o) = lo;
        };
        const uint32_t unit_base = a_off(row, q);                      // operand column of unit 4i + q: + 128 i bytes
        auto put_unit = [&](unsigned char* region, int i, float v) {
            float hi, lo;
            tc::split_tf32_trunc(v, hi, lo);
            *reinterpret_cast<float*>(region + unit_base + 128 * i) = hi;
            *reinterpret_cast<float*>(region + G_PLANE_A + unit_base + 128 * i) = lo;
        };
        auto load_x = [&](int t, float (&xv)[10]) {
#pragma unroll
            for (int i = 0; i < 10; ++i) {
                const int k = kx0 + i;
                xv[i] = (live && k < F_in) ? __ldg(p.noise + ((size_t)path * T + t) * F_in + k) : 0.f;
            }
        };
        auto store_x = [&](const float (&xv)[10]) {
#pragma unroll
            for (int i = 0; i < 10; ++i) put(AX, kx0 + i, xv[i]);
        };
        auto signal_operands = [&]() {                                 // operands written, accumulators read: the MMA thread may go on
            tc::fence_async_smem();
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&bar_a);
        };
        // one LSTM layer's epilogue: gates -> cell -> h (region Hreg) -> [LeakyReLU] LayerNorm -> a[] (registers)
        const uint32_t zcol = trow + (uint32_t)(200 * pr);             // this warp pair's 200 gate columns
        auto layer = [&](int own, float (&c)[G_UPL], const float* bp, unsigned char* Hreg, int lrelu, const float* gamma,
                         const float* beta, float (&a)[G_UPL]) {
            float hv[G_UPL];
            float ssum = 0.f;
            const float* bq = bp + 200 * pr + 4 * hh;                   // + 8 i: the four gate biases of unit 4i + q
            // gates of one unit: the lanes that hold the TMEM rows (half `own`: 0 = lanes 0-15, layer 1; 1 = lanes 16-31, layer 2)
            // keep their own four of a unit pair's eight columns and hand the other four to the other half
            // NU units at once, stage by stage: the compiler keeps a unit's ~20 dependent operations (shuffle, ex2, rcp, ...)
            // in order, so written unit by unit the chains run one after the other; written across the units they overlap.
            auto units = [&](auto nu_tag, int i0, const uint32_t* r) {
                constexpr int NU = decltype(nu_tag)::value;
                float z[NU][4];
#pragma unroll
                for (int k = 0; k < NU; ++k) {
                    const float4 b4 = *reinterpret_cast<const float4*>(bq + 8 * (i0 + k));
                    const float bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const uint32_t o = __shfl_sync(0xffffffffu, r[8 * k + 4 * (1 - own) + g], l16 + 16 * own);
                        z[k][g] = __uint_as_float(hh == own ? r[8 * k + 4 * own + g] : o) + bb[g];
                    }
                }
                // sigmoids with 5 ex2 + 3 rcp per unit: i*cand and o*sigmoid(c) share a reciprocal each
                float e[NU][4];
#pragma unroll
                for (int k = 0; k < NU; ++k)
#pragma unroll
                    for (int g = 0; g < 4; ++g) e[k][g] = ex2_fast(z[k][g] * -1.4426950408889634f);
                float rf[NU], ric[NU];
#pragma unroll
                for (int k = 0; k < NU; ++k) { rf[k] = rcp_fast(1.f + e[k][1]); ric[k] = rcp_fast((1.f + e[k][0]) * (1.f + e[k][2])); }
                float es[NU];
#pragma unroll
                for (int k = 0; k < NU; ++k) { c[i0 + k] = fmaf(rf[k], c[i0 + k], ric[k]); es[k] = ex2_fast(c[i0 + k] * -1.4426950408889634f); }
                float h[NU];
#pragma unroll
                for (int k = 0; k < NU; ++k) h[k] = rcp_fast((1.f + e[k][3]) * (1.f + es[k]));
#pragma unroll
                for (int k = 0; k < NU; ++k) {
                    put_unit(Hreg, i0 + k, h[k]);
                    hv[i0 + k] = (lrelu && !(h[k] > 0.f)) ? alpha * h[k] : h[k];
                    ssum += hv[i0 + k];
                }
            };
#pragma unroll
            for (int i0 = 0; i0 < 24; i0 += 4) {                        // four units of both lane halves per 32-column load
                uint32_t r[32];
                tc::tmem_ld32_nowait(zcol + (uint32_t)(8 * i0), r);
                tc::tmem_ld_wait();
                units(std::integral_constant<int, 4>{}, i0, r);
            }
            {
                uint32_t r[8];
                tc::tmem_ld8_nowait(zcol + 8u * 24u, r);
                tc::tmem_ld_wait();
                units(std::integral_constant<int, 1>{}, 24, r);
            }
            lnA[row * 4 + q] = ssum;
            epi_sync();
            const float mean = ((lnA[row * 4] + lnA[row * 4 + 1]) + (lnA[row * 4 + 2] + lnA[row * 4 + 3])) / (float)GU;
            float ss = 0.f;
#pragma unroll
            for (int j = 0; j < G_UPL; ++j) { const float d = hv[j] - mean; ss = fmaf(d, d, ss); }
            lnB[row * 4 + q] = ss;
            epi_sync();
            const float inv = rsqrtf(((lnB[row * 4] + lnB[row * 4 + 1]) + (lnB[row * 4 + 2] + lnB[row * 4 + 3])) / (float)GU + p.ln_eps);
#pragma unroll
            for (int j = 0; j < G_UPL; ++j) a[j] = (hv[j] - mean) * inv * gamma[G_UPL * q + j] + beta[G_UPL * q + j];
        };

        {   // x_0, then the first operands are ready
            float xv[10];
            load_x(0, xv);
            store_x(xv);
            signal_operands();
        }
        const uint32_t dcol = trow + 400u;                             // the Dense layer's accumulators: 40 columns, lanes 0-15
        for (int t = 0; t < T; ++t) {
            float xn[10];
            if (t + 1 < T) load_x(t + 1, xn);                          // (in flight under the first layer's epilogue)
            // ---- layer 1
            tc::mbar_wait<false>(&bar_z, par_z); par_z ^= 1;
            tc::tc_fence_after();
            float a[G_UPL];
            layer(0, c1, b1p, H1, p.lrelu1, lng, lng + GU, a);
#pragma unroll
            for (int j = 0; j < G_UPL; ++j) put_unit(AX, j, a[j]);
            signal_operands();                                         // A(t): a1_t, h1_t in place, layer 1's lanes read
            // ---- layer 2
            tc::mbar_wait<false>(&bar_z, par_z); par_z ^= 1;
            tc::tc_fence_after();
            layer(1, c2, b2p, H2, p.lrelu2, lng + 2 * GU, lng + 3 * GU, a);
#pragma unroll
            for (int j = 0; j < G_UPL; ++j) put_unit(AX, j, a[j]);     // layer 2's a1-half is complete: a2_t takes the region
            signal_operands();                                         // C(t): a2_t, h2_t in place, layer 2's lanes read
            // ---- Dense: out = a2 Kd + bd, one more MMA chain; the two warps of a TMEM quarter store half the columns each
            tc::mbar_wait<false>(&bar_z, par_z); par_z ^= 1;
            tc::tc_fence_after();
            {
                uint32_t r[32], r8[8];
                tc::tmem_ld32_nowait(dcol, r);
                tc::tmem_ld8_nowait(dcol + 32u, r8);
                tc::tmem_ld_wait();
                if (hh == 0 && live) {
                    float* o = p.out + ((size_t)path * T + t) * F_out;
                    if (F_out == 36) {                                 // rows of 144 bytes, 16-byte aligned
#pragma unroll
                        for (int j4 = 0; j4 < 9; ++j4) {
                            if ((j4 < 5) != (pr == 0)) continue;
                            const float4 b4 = __ldg(reinterpret_cast<const float4*>(bdg) + j4);
                            const uint32_t* v = j4 < 8 ? r + 4 * j4 : r8;
                            reinterpret_cast<float4*>(o)[j4] = make_float4(__uint_as_float(v[0]) + b4.x, __uint_as_float(v[1]) + b4.y,
                                                                           __uint_as_float(v[2]) + b4.z, __uint_as_float(v[3]) + b4.w);
                        }
                    } else if (pr == 0) {
#pragma unroll
                        for (int j = 0; j < 36; ++j)
                            if (j < F_out) o[j] = __uint_as_float(j < 32 ? r[j] : r8[j - 32]) + __ldg(bdg + j);
                    }
                }
            }
            if (t + 1 < T) {                                           // the Dense chain is complete: x_{t+1} takes the region
                store_x(xn);
                signal_operands();                                     // B(t): x_{t+1} in place, the Dense lanes read
            }
        }
        tc::tc_fence_before();
    } else if (warp == G_EPI / 32) {
        // ============================================================ MMA issue
        if (lane == 0) {
            const uint32_t i0 = tc::idesc_tf32(64, TCN0), i1 = tc::idesc_tf32(64, TCN1), id = tc::idesc_tf32(64, G_ND);
            const uint32_t h1a = tc::smem_u32(H1), axa = tc::smem_u32(AX), h2a = tc::smem_u32(H2), rga = tc::smem_u32(ring);
            unsigned par_a = 0;
            uint32_t g = 0;                                            // K-steps issued so far: stage = g & 1
            // one K = 8 step: three MMA pairs (3xTF32) on the stage the loader filled, then the stage is handed back
            auto kstep = [&](uint32_t abase, uint32_t d, bool first) {
                const uint32_t st = g & 1;
                tc::mbar_wait<false>(&bar_full[st], (g >> 1) & 1);
                tc::tc_fence_after();
                const uint32_t bbase = rga + st * G_STAGE;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const uint32_t aa = abase + (c == 1 ? G_PLANE_A : 0u), bb = bbase + (c == 2 ? G_PLANE_B : 0u);
                    const uint64_t ad = tc::smem_desc(aa, G_LBO, G_SBO_A);
                    tc::mma_tf32(d, ad, tc::smem_desc(bb, G_LBO, G_SBO_B), i0, !first || c > 0);
                    tc::mma_tf32(d + TCN0, ad, tc::smem_desc(bb + (TCN0 / 8) * G_SBO_B, G_LBO, G_SBO_B), i1, !first || c > 0);
                }
                tc::mma_commit(&bar_empty[st]);                         // the stage is free once these MMAs have read it
                ++g;
            };
            auto kstep_dense = [&](uint32_t abase, uint32_t d, bool first) {
                const uint32_t st = g & 1;
                tc::mbar_wait<false>(&bar_full[st], (g >> 1) & 1);
                tc::tc_fence_after();
                const uint32_t bbase = rga + st * G_STAGE;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const uint32_t aa = abase + (c == 1 ? G_PLANE_A : 0u), bb = bbase + (c == 2 ? G_PLANE_D : 0u);
                    tc::mma_tf32(d, tc::smem_desc(aa, G_LBO, G_SBO_A), tc::smem_desc(bb, G_LBO, G_SBO_B), id, !first || c > 0);
                }
                tc::mma_commit(&bar_empty[st]);
                ++g;
            };
            const uint32_t z1 = tmem, z2 = tmem + (16u << 16), zd = tmem + 400u;   // layer 2: lanes 16-31 of every TMEM quarter
            // Issue order (the loader streams the weight blocks in the same order).  Every chain starts with the half whose
            // operand is ready first, so it runs under the epilogue that produces the other half's operand:
            //   h1_{t-1} U1 | per step:  x_t W1 -> z1 ; h2_{t-1} U2 ; [A(t)] a1_t W2 -> z2 ; h1_t U1 (next step's) ; [C(t)] a2_t Kd -> zd ; [B(t)]
            tc::mbar_wait<false>(&bar_a, par_a); par_a ^= 1;                   // x_0 and the zero states
            tc::tc_fence_after();
            for (int ks = 0; ks < GKH / 8; ++ks) kstep(h1a + (uint32_t)ks * 256u, z1, ks == 0);
            for (int t = 0; t < T; ++t) {
                if (t > 0) { tc::mbar_wait<false>(&bar_a, par_a); par_a ^= 1; tc::tc_fence_after(); }   // B(t-1)
                for (int ks = 0; ks < GKX / 8; ++ks) kstep(axa + (uint32_t)ks * 256u, z1, false);
                tc::mma_commit(&bar_z);                                 // layer 1's accumulators are complete
                for (int ks = 0; ks < GKH / 8; ++ks) kstep(h2a + (uint32_t)ks * 256u, z2, ks == 0);
                tc::mbar_wait<false>(&bar_a, par_a); par_a ^= 1;               // A(t)
                tc::tc_fence_after();
                for (int ks = 0; ks < GKH / 8; ++ks) kstep(axa + (uint32_t)ks * 256u, z2, false);
                tc::mma_commit(&bar_z);                                 // layer 2's accumulators are complete
                if (t + 1 < T)
                    for (int ks = 0; ks < GKH / 8; ++ks) kstep(h1a + (uint32_t)ks * 256u, z1, ks == 0);
                tc::mbar_wait<false>(&bar_a, par_a); par_a ^= 1;               // C(t)
                tc::tc_fence_after();
                for (int ks = 0; ks < G_NKD; ++ks) kstep_dense(axa + (uint32_t)ks * 256u, zd, ks == 0);
                tc::mma_commit(&bar_z);                                 // the Dense accumulators are complete
            }
        }
        __syncwarp();
    } else {
        // ============================================================ weight loader (same order as the MMA issue)
        if (lane == 0) {
            uint32_t g = 0;
            auto load = [&](const unsigned char* src, uint32_t bytes) {
                const uint32_t st = g & 1;
                tc::mbar_wait(&bar_empty[st], ((g >> 1) & 1) ^ 1);
                tc::mbar_expect_tx(&bar_full[st], bytes);
                tc::bulk_g2s(ring + st * G_STAGE, src, bytes, &bar_full[st]);
                ++g;
            };
            const unsigned char* bx = p.tcw;                                   // 5 blocks: W1 rows of x
            const unsigned char* bh1 = bx + (size_t)(GKX / 8) * G_STAGE;       // 13: U1
            const unsigned char* bh2 = bh1 + (size_t)(GKH / 8) * G_STAGE;      // 13: U2
            const unsigned char* ba1 = bh2 + (size_t)(GKH / 8) * G_STAGE;      // 13: W2
            const unsigned char* bkd = ba1 + (size_t)(GKH / 8) * G_STAGE;      // 13 small: Kd
            for (int ks = 0; ks < GKH / 8; ++ks) load(bh1 + (size_t)ks * G_STAGE, G_STAGE);
            for (int t = 0; t < T; ++t) {
                for (int ks = 0; ks < GKX / 8; ++ks) load(bx + (size_t)ks * G_STAGE, G_STAGE);
                for (int ks = 0; ks < GKH / 8; ++ks) load(bh2 + (size_t)ks * G_STAGE, G_STAGE);
                for (int ks = 0; ks < GKH / 8; ++ks) load(ba1 + (size_t)ks * G_STAGE, G_STAGE);
                if (t + 1 < T)
                    for (int ks = 0; ks < GKH / 8; ++ks) load(bh1 + (size_t)ks * G_STAGE, G_STAGE);
                for (int ks = 0; ks < G_NKD; ++ks) load(bkd + (size_t)ks * G_BLOCK_D, G_BLOCK_D);
            }
        }
        __syncwarp();
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) tc::tmem_free<512>(tmem);
}

// B blocks of the two LSTM layers for generator_tc_kernel: block b = K-step b of a time step in issue order (18 of layer 1: x
// rows 0..39, then h1 rows 0..103; 26 of layer 2: h2 rows first, then a1 rows; rows past the real input sizes are zero), each [hi | lo] plane
// 400 gate columns x 8 K in the canonical K-major layout, gate column 4u+g <- Keras column g*U+u.
__global__ void generator_tc_prepare_kernel(const float* W, int F_in, int F_out, unsigned char* tcw) {
    const int U = GU, G = GG;
    const float* W1 = W;
    const float* U1 = W1 + (size_t)F_in * G;
    const float* W2 = U1 + (size_t)U * G + G + 2 * U;
    const float* U2 = W2 + (size_t)U * G;
    const float* Kd = U2 + (size_t)U * G + G + 2 * U;
    const int total = G_NKS * 8 * G;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int n = i % G, kk = (i / G) % 8, b = i / (8 * G);
        const int u = gate_col_unit(n), g = gate_col_gate(n), col = g * U + u;
        float v = 0.f;
        if (b < GKX / 8) {                                             // layer 1: x rows, then h1 rows
            const int k = b * 8 + kk;
            if (k < F_in) v = W1[(size_t)k * G + col];
        } else if (b < G_NKS1) {
            const int k = (b - GKX / 8) * 8 + kk;
            if (k < U) v = U1[(size_t)k * G + col];
        } else if (b < G_NKS1 + GKH / 8) {                             // layer 2: h2 rows (recurrent half), then a1 rows
            const int k = (b - G_NKS1) * 8 + kk;
            if (k < U) v = U2[(size_t)k * G + col];
        } else {
            const int k = (b - G_NKS1 - GKH / 8) * 8 + kk;
            if (k < U) v = W2[(size_t)k * G + col];
        }
        float hi, lo;
        tc::split_tf32(v, hi, lo);
        const size_t o = (size_t)b * G_STAGE + (size_t)(n >> 3) * G_SBO_B + (size_t)(kk >> 2) * 128 + (size_t)(n & 7) * 16 + (size_t)(kk & 3) * 4;
        *reinterpret_cast<float*>(tcw + o) = hi;
        *reinterpret_cast<float*>(tcw + G_PLANE_B + o) = lo;
    }
    // the Dense kernel: 13 blocks of [hi | lo] planes, 40 output columns (36 + zero padding) x 8 K
    const int total_d = G_NKD * 8 * G_ND;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total_d; i += gridDim.x * blockDim.x) {
        const int n = i % G_ND, kk = (i / G_ND) % 8, b = i / (8 * G_ND);
        const int k = b * 8 + kk;
        const float v = (k < U && n < F_out) ? Kd[(size_t)k * F_out + n] : 0.f;
        float hi, lo;
        tc::split_tf32(v, hi, lo);
        const size_t o = (size_t)G_NKS * G_STAGE + (size_t)b * G_BLOCK_D + (size_t)(n >> 3) * G_SBO_B + (size_t)(kk >> 2) * 128 +
                         (size_t)(n & 7) * 16 + (size_t)(kk & 3) * 4;
        *reinterpret_cast<float*>(tcw + o) = hi;
        *reinterpret_cast<float*>(tcw + G_PLANE_D + o) = lo;
    }
}

}  // namespace hfrep

using namespace hfrep;

extern "C" int hfrep_tc_selftest(const float* A, const float* B, int K, float* D, int lane_half, void* stream) {
    HFREP_REQUIRE(A && B && D, "tc_selftest: null argument");
    HFREP_REQUIRE(K >= 8 && K % 8 == 0 && K <= 48, "tc_selftest: K=%d (a multiple of 8 up to 48)", K);
    const size_t smem = (size_t)2 * (TCM + TCN) * K * 4;
    HFREP_CUDA_OK(allow_max_dynamic_smem(tc_selftest_kernel));
    tc_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(A, B, K, D, lane_half ? 1 : 0);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

extern "C" size_t hfrep_generator_tc_weight_bytes(int F_in, int units, int F_out) {
    if (units != GU || F_in < 1 || F_in > GKX || F_out < 1 || F_out > 36) return 0;       // shapes the tensor-core kernel takes
    return (size_t)G_NKS * G_STAGE + (size_t)G_NKD * G_BLOCK_D;
}

extern "C" int hfrep_generator_tc_prepare(const float* weights, int F_in, int units, int F_out, void* tc_weights, void* stream) {
    HFREP_REQUIRE(weights && tc_weights, "generator_tc_prepare: null argument");
    HFREP_REQUIRE(hfrep_generator_tc_weight_bytes(F_in, units, F_out) != 0, "generator_tc_prepare: F_in=%d units=%d F_out=%d (units = %d, F_in <= %d, F_out <= 36)",
                  F_in, units, F_out, GU, GKX);
    generator_tc_prepare_kernel<<<148, 256, 0, (cudaStream_t)stream>>>(weights, F_in, F_out, (unsigned char*)tc_weights);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

extern "C" int hfrep_generator_forward_tc(const float* weights, const void* tc_weights, int F_in, int units, int F_out,
                                          int lrelu_after_lstm1, int lrelu_after_lstm2, float alpha, float ln_eps,
                                          const float* noise, int n_paths, int T, float* out, void* stream) {
    HFREP_REQUIRE(weights && tc_weights && noise && out, "generator_forward_tc: null argument");
    HFREP_REQUIRE(hfrep_generator_tc_weight_bytes(F_in, units, F_out) != 0, "generator_forward_tc: F_in=%d units=%d F_out=%d not taken", F_in, units, F_out);
    HFREP_REQUIRE(n_paths >= 0 && T >= 1, "generator_forward_tc: n_paths=%d T=%d", n_paths, T);
    if (n_paths == 0) return HFREP_OK;
    GenTcParams p;
    p.W = weights; p.tcw = (const unsigned char*)tc_weights; p.noise = noise; p.out = out; p.n = n_paths; p.T = T;
    p.F_in = F_in; p.F_out = F_out; p.lrelu1 = lrelu_after_lstm1; p.lrelu2 = lrelu_after_lstm2; p.alpha = alpha; p.ln_eps = ln_eps;
    const size_t smem = 6 * (size_t)G_PLANE_A + 2 * (size_t)G_STAGE + sizeof(float) * (2 * GG + 4 * GU + 512);
    HFREP_CUDA_OK(allow_max_dynamic_smem(generator_tc_kernel));
    generator_tc_kernel<<<(n_paths + 63) / 64, 320, smem, (cudaStream_t)stream>>>(p);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

extern "C" int hfrep_tc_mma_rate(int N, int iters, void* stream) {
    HFREP_REQUIRE(N >= 8 && N <= 256 && N % 8 == 0 && iters >= 1, "tc_mma_rate: N=%d iters=%d", N, iters);
    tc_mma_rate_kernel<<<sm_count(), 128, (64 + 256) * 8 * 4, (cudaStream_t)stream>>>(N, iters);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}
