/*
 * hfrep.h — C ABI of the B200-native replication hot path (libhfrep.so, sm_100a).
 *
 * This is the drop-in boundary underneath the Python entry points of the reference's
 * Autoencoder_encapsulate.py / helper.py (the reference has no FFI of its own; its
 * "interface" is those Python signatures).  Every entry point below names
 * the reference lines whose arithmetic it replaces.  Plain C: caller-owned buffers, sizes
 * and a CUDA stream handle (cudaStream_t passed as void*); no torch / C++ types.
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the parameter name ends in _host;
 *   - matrices are row-major; "f32"/"f64" = IEEE float/double; indices are int32;
 *   - every call is asynchronous on `stream` and allocates nothing: buffers, scratch included,
 *     are passed in by the caller (one exception: hfrep_fit_windows_static takes its scratch
 *     from an array built into the library instead of from the caller);
 *   - return value 0 = ok, <0 = error; hfrep_last_error() returns a thread-local message;
 *   - no CPU fallback exists: without a CUDA device the calls return HFREP_ERR_CUDA.
 */
#ifndef HFREP_H
#define HFREP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HFREP_OK 0
#define HFREP_ERR_ARG (-1)    /* bad argument / unsupported shape */
#define HFREP_ERR_CUDA (-2)   /* CUDA runtime error (message has the CUDA string) */

#define HFREP_MAX_E 24        /* ETF width handled by the narrow kernels (reference: 22) */
#define HFREP_MAX_L 24        /* latent width (reference sweep: 1..21) */
#define HFREP_MAX_F 32        /* funds per evaluation (reference: 13) */
#define HFREP_N_STATS 11      /* statistics per return series, see hfrep_series_metrics */

const char* hfrep_last_error(void);
int hfrep_version(void);

/* ---------------------------------------------------------------------------------------
 * Window index tables (host side, bit-exact int32).
 *   kind 0 rolling   : start=i, len=window, i=0..T-window           (Autoencoder_encapsulate.py:145-156; count T-window)
 *   kind 1 expanding : start=0, len=i,      i=2..T-1               (Autoencoder_encapsulate.py:117,126; count T-2)
 *   kind 2 expost    : start=i, len=window, i=1..T-window-1         (helper.py:120-121; count T-window-1)
 * Writes up to `cap` entries to starts_host/lens_host and returns the count (or <0).
 * Random windows (helper.py:55-58) come from python's `random.randint`; the Python facade draws
 * them and passes the table in — there is nothing to compute natively.
 */
int hfrep_window_index(int kind, int T, int window, int32_t* starts_host, int32_t* lens_host, int cap);

/* ---------------------------------------------------------------------------------------
 * Kernel (a): window gather.  out[w][r][c] = R[starts[w]+r][col_lo+c], r<len, c<ncols, for
 * n windows of equal length `len`; replaces the python slicing loops of helper.random_sampling
 * (helper.py:55-58), helper.factor_hf_split (helper.py:145-152) and the window slices of
 * Autoencoder_encapsulate.py:149-150.  If `minmax` != 0 every window is MinMax-scaled per
 * column on the fly exactly as MinMaxScaler().fit_transform does (Autoencoder_encapsulate.py:62-65):
 * (x-min)*(1/(max-min)) in float64, zero range -> scale 1, rounded to f32.
 *   R   f32 [T][C]      starts int32 [n]      out f32 [n][len][ncols]
 */
int hfrep_window_gather(const float* R, int T, int C, const int32_t* starts, int n, int len,
                        int col_lo, int ncols, int minmax, float* out, void* stream);
/* The same from a float64 returns matrix (the reference's DataFrames are float64): min, max and the scaling are taken on
 * the float64 values and the result is rounded to f32 once — MinMaxScaler on float64 (Autoencoder_encapsulate.py:62-65)
 * followed by Keras' cast to float32 (:83), bit for bit.  Without minmax the values are rounded to f32. */
int hfrep_window_gather_f64(const double* R, int T, int C, const int32_t* starts, int n, int len,
                            int col_lo, int ncols, int minmax, float* out, void* stream);

/* Rolling ex-post volatility (helper.py:121 + :75/:87): sigma[i][e] = sqrt(var(R[start_i..start_i+window), e], ddof=1)),
 * f64 in / f64 out.  R f64 [T][E], starts int32 [n], sigma f64 [n][E]. */
int hfrep_rolling_sigma(const double* R, int T, int E, const int32_t* starts, int n, int window,
                        double* sigma, void* stream);

/* ---------------------------------------------------------------------------------------
 * Kernel (b): batched autoencoder fit.  Replaces Autoencoder (Autoencoder_encapsulate.py:19-35)
 * + AE.train (Autoencoder_encapsulate.py:72-96), i.e. Keras Dense(no bias)/LeakyReLU/MSE/Nadam/
 * fit(validation_split, batch_size)/EarlyStopping, for n_tasks independent tasks in one launch.
 */
typedef struct hfrep_fit_cfg {
    int32_t E;            /* ETF columns (22), <= HFREP_MAX_E */
    int32_t L_max;        /* largest latent width among the tasks of this call, <= HFREP_MAX_L */
    int32_t T;            /* rows per training block (168) */
    int32_t n_fit;        /* floor(T*(1-validation_split)) = 126: rows [0,n_fit) train, [n_fit,T) validate */
    int32_t batch;        /* 48 */
    int32_t max_epochs;   /* 1000 */
    int32_t patience;     /* 5 */
    int32_t early_stop;   /* 1 = EarlyStopping(val_loss, patience), 0 = run max_epochs */
    float alpha;          /* LeakyReLU slope .2 */
    float lr, beta_1, beta_2, epsilon, schedule_decay, decay_base;  /* Nadam(): 1e-3,.9,.999,1e-7,.004,.96 */
    int32_t n_x;          /* number of training blocks in X; 1 = all tasks share block 0 */
    int32_t warps_per_task; /* 0 = choose from this call's n_tasks; 1, 2, 3 = throughput mappings (1 needs many tasks
                               per SM); 6 = latency mapping for a handful of tasks (one CTA per task);
                               HFREP_WPT_CLUSTER = a cluster of six CTAs (six SMs) per task, for so few tasks that every
                               cluster has SMs of its own (falls back to 6 for shapes it does not take) */
    int32_t dump_step;    /* global step (0-based) whose gradients go to grad_out, if grad_out != NULL */
} hfrep_fit_cfg;

#define HFREP_WPT_CLUSTER 60

void hfrep_fit_cfg_default(hfrep_fit_cfg* cfg);

/* Nadam's scalar schedule is task-independent: fill sched f32 [n_steps][HFREP_SCHED_STRIDE] (host) with
 * {1-m_sched_t, 1-m_sched_{t+1}, 1-beta_2^t, 1-mu_t, mu_{t+1}, mu_t, 0, 0}, t=1..n_steps (Keras 2.7
 * Nadam._prepare_local in float32, powers correctly rounded).  The first three DIVIDE the gradient and the two moments:
 * the kernels apply Keras' operation order with IEEE division and square root, every operation rounded on its own. */
#define HFREP_SCHED_STRIDE 8
int hfrep_nadam_schedule(const hfrep_fit_cfg* cfg, int n_steps, float* sched_host);

/*   X        f32 [n_x][T][E]        training blocks, already MinMax-scaled
 *   task_L   int32 [n_tasks]        latent width of each task (<= L_max)
 *   task_x   int32 [n_tasks]|NULL   block of each task (NULL: block 0)
 *   W0       f32 [n_tasks][w_stride] initial kernels: We (E x L) then Wd (L x E), packed with the task's own L
 *   perm     int32 [.. ][epochs][n_fit] per-epoch row order; NULL = no shuffling; perm_task_stride
 *            (in int32 elements) = 0 shares one table between all tasks
 *   task_seed uint32 [n_tasks] | NULL  device-side epoch shuffles (Keras fit's shuffle=True without a host table): when
 *            perm is NULL and task_seed is not, every epoch's row order is the stable argsort of the hash keys
 *            fmix32(seed*0x9E3779B1 ^ (epoch+1)*0x85EBCA77 ^ (row+1)*0xC2B2AE3D) of the task's seed (murmur3 finaliser)
 *   sched    f32 [>= max_epochs*ceil(n_fit/batch)][HFREP_SCHED_STRIDE]  from hfrep_nadam_schedule, on the device, 16-byte aligned
 *   W_out    f32 [n_tasks][w_stride]       final kernels, same packing (floats past 2*E*L of a row are written as zero)
 *   loss_hist/val_hist f32 [n_tasks][max_epochs]   history.history['loss'|'val_loss']; entries past epochs_run are
 *            written as NaN — no output buffer needs initialising
 *   epochs_run int32 [n_tasks];  status int32 [n_tasks] (0 ok, 1 non-finite loss)
 *   grad_out f32 [n_tasks][w_stride] | NULL   gradients (dWe, dWd) of step cfg->dump_step
 * Every sum that feeds the weights follows one fixed order in every mapping (fused chains over the ETFs / latent
 * units, six interleaved chains over the batch rows: DESIGN.md §5), so W_out does not depend on warps_per_task, on the
 * task's place in the batch or on the staging mode, and equals the CPU oracle's canonical-order result bit for bit.
 * With early_stop = 1 and no validation rows (n_fit == T) there is no val_loss to monitor: like Keras, the run
 * trains all max_epochs.
 */
int hfrep_fit_batched(const hfrep_fit_cfg* cfg, const float* X, int n_tasks, const int32_t* task_L,
                      const int32_t* task_x, const float* W0, int w_stride, const int32_t* perm,
                      int64_t perm_task_stride, const uint32_t* task_seed, const float* sched, float* W_out,
                      float* loss_hist, float* val_hist, int32_t* epochs_run, int32_t* status, float* grad_out,
                      void* stream);

/* Rolling refit (the north_star's per-window task; SURVEY §8d "Mode W", not in the reference): task i trains on the
 * window R[win_starts[task_win[i]] .. +cfg->T)[col_lo .. col_lo+cfg->E), MinMax-scaled per window exactly as
 * AE.__init__ scales its training half (Autoencoder_encapsulate.py:62-65).  Kernel (a) writes the scaled blocks
 * into the caller's workspace `ws` (f32 [n_windows][cfg->T][cfg->E], size from hfrep_fit_windows_workspace, 16-byte
 * aligned), kernel (b) fits on them.  Other arguments as hfrep_fit_batched. */
size_t hfrep_fit_windows_workspace(int n_windows, int T_win, int E);
int hfrep_fit_windows(const hfrep_fit_cfg* cfg, const float* R, int T_all, int C, int col_lo,
                      const int32_t* win_starts, int n_windows, int n_tasks, const int32_t* task_L,
                      const int32_t* task_win, const float* W0, int w_stride, const int32_t* perm,
                      int64_t perm_task_stride, const uint32_t* task_seed, const float* sched, void* ws,
                      size_t ws_bytes, float* W_out, float* loss_hist, float* val_hist, int32_t* epochs_run,
                      int32_t* status, void* stream);

/* hfrep_fit_windows for callers that cannot allocate device memory.  (No counterpart in the reference: the reference
 * never touches a GPU.)  Same arguments as hfrep_fit_windows minus the workspace: the scaled window blocks go into a
 * fixed array built into the library (HFREP_STATIC_SCRATCH_BYTES per GPU, reserved from load to exit).  Meant for
 * hosts whose own arena already owns the GPU's memory, or that cannot take the device-wide stall of an allocation
 * while other streams are fitting.  Limits: n_windows * cfg->T * cfg->E * 4 bytes must fit the array, else the call
 * returns HFREP_ERR_ARG and names both sizes.  The first call on each device looks up the array's device address; make
 * that call outside stream capture, and before the host's arena takes the rest of the GPU (CUDA loads the array and
 * each kernel variant lazily, at first use, unless CUDA_MODULE_LOADING=EAGER).  All calls on a device share the array — calls that overlap on the GPU race on
 * it; calls ordered on one stream (or by events) are safe.  Nothing needs clearing between calls. */
#define HFREP_STATIC_SCRATCH_BYTES (4u << 20)
int hfrep_fit_windows_static(const hfrep_fit_cfg* cfg, const float* R, int T_all, int C, int col_lo,
                             const int32_t* win_starts, int n_windows, int n_tasks, const int32_t* task_L,
                             const int32_t* task_win, const float* W0, int w_stride, const int32_t* perm,
                             int64_t perm_task_stride, const uint32_t* task_seed, const float* sched, float* W_out,
                             float* loss_hist, float* val_hist, int32_t* epochs_run, int32_t* status, void* stream);

/* ---------------------------------------------------------------------------------------
 * Kernel (c): fused out-of-sample predict + rolling OLS + replica + ex-post + metrics, one
 * task (= one trained autoencoder) per CTA.  Replaces AE.model_IS_r2/RMSE (:107-113),
 * AE.model_OOS_r2/RMSE (:115-131), AE.ante (:133-201) with helper.normalization (helper.py:10-17),
 * AE.post (:203-208) with helper.ex_post_return/transaction_cost/price_impact (helper.py:65-92,112-131),
 * AE.turnover (:210-224), annualized_sharpe_ratio (autoencoder_v4.ipynb:715-721) and the two
 * metrics the reference does not have (tracking error, correlation).
 */
typedef struct hfrep_eval_cfg {
    int32_t E, F;         /* ETF columns, fund columns (<= HFREP_MAX_F) */
    int32_t T_train;      /* rows of Xtr (168) */
    int32_t T_test;       /* rows of the test block (169) */
    int32_t window;       /* OLS / ex-post window (24), <= 64 */
    int32_t beta_rolling; /* 0 = reference behaviour (window 0's beta/norm for every month, :167); 1 = month k uses
                             window k's (the evident intent) — needs the `weights` output as storage */
    int32_t want_oos;     /* 1 = also the expanding-prefix quality loops (:115-131) */
    float alpha;          /* LeakyReLU slope */
    double cost_param;    /* .05  (helper.py:65,83) */
    double phi;           /* .5   (helper.py:83) */
} hfrep_eval_cfg;

void hfrep_eval_cfg_default(hfrep_eval_cfg* cfg);

/* n_oos = T_test - window - 1 replica months (144); n_pre = T_test - 2 expanding prefixes (167).
 *   Xtr    f64 [T_train][E]  MinMax-scaled train block (as AE.__init__ stores it, :65)
 *   Xte    f64 [T_test][E]   raw test ETF rows (:67);  Yte f64 [T_test][F] test fund rows
 *   rf     f64 [T_test]      risk-free rows aligned with the test block
 *   sigma  f64 [n_oos-1][E]  from hfrep_rolling_sigma over the last n_oos+window ETF rows, starts 1..n_oos-1
 *   W      f32 [n_tasks][w_stride], task_L int32 [n_tasks]: trained kernels from hfrep_fit_batched
 *   task_row0 int32 [n_tasks] | NULL: rolling refit (one autoencoder per window): task i evaluates rows
 *          [task_row0[i], task_row0[i]+T_test) of Xte/Yte/rf, which then hold the whole history (NULL: row 0)
 * outputs (any may be NULL to skip the store):
 *   ante, post f64 [n_tasks][n_oos][F];  turnover f64 [n_tasks][F]
 *   metrics f64 [n_tasks][2][HFREP_N_STATS][F]: the statistics of hfrep_series_metrics for ante, then for post
 *          (left untouched when fewer than 4 replica months exist)
 *   quality f64 [n_tasks][2 + 2*n_pre]: IS r2, IS rmse, OOS r2[n_pre], OOS rmse[n_pre]
 *   mask u8 [n_tasks][n_oos][E]  (1 = LeakyReLU slope applied, i.e. pre-activation < 0, :164-166)
 *   weights f64 [n_tasks][n_oos][E][F]  strat_weight_on_etf (:167); large — facade use only
 *   beta f64 [n_tasks][L_stride=HFREP_MAX_L][F] window-0 OLS coefficients; norm f64 [n_tasks][F]
 *   status int32 [n_tasks] (0 ok, 2 rank-deficient OLS window)
 */
int hfrep_predict_metrics(const hfrep_eval_cfg* cfg, const double* Xtr, const double* Xte, const double* Yte,
                          const double* rf, const double* sigma, int n_tasks, const int32_t* task_L,
                          const int32_t* task_row0, const float* W, int w_stride, double* ante, double* post, double* turnover,
                          double* metrics, double* quality, uint8_t* mask, double* weights, double* beta,
                          double* norm, int32_t* status, void* stream);

/* Network forward pass for arbitrary rows (Autoencoder.call :32-35, Model.predict, encoder.predict :140):
 * x f32 [n][E]; W f32 packed We (E x L), Wd (L x E); encode_only != 0 -> out f32 [n][L] = lrelu(x We),
 * else out f32 [n][E] = lrelu(lrelu(x We) Wd). */
int hfrep_forward(const float* W, int E, int L, const float* x, int n, int encode_only, float alpha, float* out,
                  void* stream);

/* ---------------------------------------------------------------------------------------
 * Wide variant (BASELINE configs[3]: 500 ETFs x 10 000 funds, rolling refit) of kernels (b) and (c): same arithmetic for
 * shapes whose per-task state does not fit shared memory, one CTA per task.  E <= 512 and L <= 32 run a register-resident
 * kernel (ETF columns dealt to threads, weights in registers, Nadam state in the workspace); other shapes a general kernel
 * with weights and Nadam state in the workspace.
 *   hfrep_fit_wide: X f32 [n_x][T][E] scaled blocks (e.g. the workspace of hfrep_window_gather with minmax=1), every task
 *   the same latent width L <= 64; W0 / W_out f32 [n_tasks][2*E*L] (We then Wd); ws f32, hfrep_fit_wide_workspace bytes
 *   (at least 6*E*L floats per task);
 *   cfg supplies E (<= 4096), T, n_fit, batch, epochs, patience, Nadam constants, n_x.
 *   hfrep_predict_wide: task i evaluates rows [task_row0[i], +window+2) of the full histories etf f64 [T][E], hfd f64 [T][F],
 *   rf f64 [T]: latent factors of those rows (float32), OLS of every fund on the first `window` rows (float64 QR; for
 *   L > window the minimum-norm solution pinv gives, via the QR of the transposed block; window <= 64), normalisation, LeakyReLU mask from row `window`, and replica[i][f] = the replica return of row
 *   window+1 (Autoencoder_encapsulate.py:133-193 with its own alignment; sums over ETFs regrouped through the L factors). */
size_t hfrep_fit_wide_workspace(int n_tasks, int E, int L);
int hfrep_fit_wide(const hfrep_fit_cfg* cfg, const float* X, int n_tasks, int L, const int32_t* task_x, const float* W0,
                   const int32_t* perm, int64_t perm_task_stride, const float* sched, void* ws, size_t ws_bytes,
                   float* W_out, float* loss_hist, float* val_hist, int32_t* epochs_run, int32_t* status, void* stream);
int hfrep_predict_wide(const double* etf, const double* hfd, const double* rf, int T, int E, int F, int L, int window,
                       float alpha, int n_tasks, const int32_t* task_row0, const float* W, double* replica,
                       int32_t* status, void* stream);

/* Stand-alone statistics of any (n, F) return block against `fund` (n, F) and rf (n): out f64 [HFREP_N_STATS][F], rows
 *   0 annualised Sharpe (autoencoder_v4.ipynb:715-721)      1 tracking error, 2 correlation (not in the reference; App. A.9)
 *   3 Omega(0 %), 4 Omega(10 %) (ipynb:695-705)              5 historical cVaR(5 %) (ipynb:729-757)
 *   6-8 CEQ(gamma = 2, 5, 10) (ipynb:759-770)                9 skewness, 10 excess kurtosis (pandas skew()/kurt(), ipynb:836)
 * i.e. the real_data=False columns of the notebook's data_analysis (ipynb:774-849) plus the two additive metrics. n >= 4. */
int hfrep_series_metrics(const double* series, const double* fund, const double* rf, int n, int F,
                         double* out, void* stream);

/* helper.ex_post_return (helper.py:112-131) for arbitrary weights: post[0][f] = ante[0][f];
 * post[t][f] = ante[t][f] + sum_e(transaction_cost + price_impact)(w[f][t-1], w[f][t], sigma[t-1]*param), t >= 1.
 *   ante f64 [T][F]; weights f64 [F][T][E] (the reshape_cab layout, helper.py:94-110);
 *   sigma f64 [T-1][E] from hfrep_rolling_sigma (starts 1..T-1 over the T+window ETF rows); post f64 [T][F]. */
int hfrep_ex_post_return(const double* ante, const double* weights, const double* sigma, int T, int F, int E,
                         double param, double phi, double* post, void* stream);

/* Generator inference: the forward pass of the reference's saved LSTM generator — GAN/MTSS_WGAN_GP.py:221-235
 * `build_generator`, the Keras model the notebook loads and runs at autoencoder_v4.ipynb:1277-1291 — fused in one kernel:
 * LSTM(units; sigmoid activation and recurrent activation; return_sequences) [-> LeakyReLU(alpha)] -> LayerNormalization(ln_eps)
 * -> LSTM [-> LeakyReLU] -> LayerNormalization -> Dense(F_out).  Keras 2.7 semantics (gate order i, f, c, o; biased variance).
 *   weights f32 [hfrep_generator_weight_count(F_in, units, F_out)], packed in this order:
 *     W1 [F_in][4u], U1 [u][4u], b1 [4u], gamma1 [u], beta1 [u], W2 [u][4u], U2 [u][4u], b2 [4u], gamma2 [u], beta2 [u],
 *     K [u][F_out], bias [F_out]                      (Keras kernel layout: in x out)
 *   noise f32 [n_paths][T][F_in]  ->  out f32 [n_paths][T][F_out].   F_in, units, F_out <= 128. */
size_t hfrep_generator_weight_count(int F_in, int units, int F_out);
int hfrep_generator_forward(const float* weights, int F_in, int units, int F_out, int lrelu_after_lstm1,
                            int lrelu_after_lstm2, float alpha, float ln_eps, const float* noise, int n_paths, int T,
                            float* out, void* stream);

/* The same forward pass on the tensor cores (csrc/gen_tc.cu: tcgen05 kind::tf32 MMAs in 3xTF32 — every product formed as
 * a_hi b_hi + a_lo b_hi + a_hi b_lo with FP32 accumulation in tensor memory; gate activations with ex2.approx / rcp.approx;
 * results agree with hfrep_generator_forward to about 1e-6 of the output range per step) for the reference's generator
 * shape: units = 100, F_in <= 40, F_out <= 36.  hfrep_generator_tc_weight_bytes returns 0 for any other shape.
 * hfrep_generator_tc_prepare lays the LSTM kernels out once (TF32 hi / lo planes per K = 8 step, gate columns permuted)
 * into tc_weights (device, hfrep_generator_tc_weight_bytes bytes, 16-byte aligned); hfrep_generator_forward_tc takes both the
 * packed weights (biases, LayerNorm, Dense) and tc_weights.  One CTA per 64 paths. */
size_t hfrep_generator_tc_weight_bytes(int F_in, int units, int F_out);
int hfrep_generator_tc_prepare(const float* weights, int F_in, int units, int F_out, void* tc_weights, void* stream);
int hfrep_generator_forward_tc(const float* weights, const void* tc_weights, int F_in, int units, int F_out,
                               int lrelu_after_lstm1, int lrelu_after_lstm2, float alpha, float ln_eps, const float* noise,
                               int n_paths, int T, float* out, void* stream);

/* tcgen05 self-test (the building blocks of the tensor-core generator path, csrc/tc.cuh): D f32 [64][400] =
 * A f32 [64][K] * B f32 [400][K]^T in 3xTF32 on one CTA — operands split into TF32 hi/lo pairs and written in the canonical
 * un-swizzled K-major shared-memory layout, kind::tf32 MMAs (M = 64, N = 256 + 144), FP32 accumulators in tensor memory.
 * K a multiple of 8, <= 48.  lane_half = 1 places the accumulator in lanes 16-31 of each TMEM quarter instead of lanes 0-15.
 * All pointers device memory. */
int hfrep_tc_selftest(const float* A, const float* B, int K, float* D, int lane_half, void* stream);
/* MMA issue-rate probe: one CTA per SM, each issues `iters` kind::tf32 MMAs of M = 64, N (a multiple of 8, <= 256), K = 8 on
 * zeroed operands; time it with CUDA events on `stream`. */
int hfrep_tc_mma_rate(int N, int iters, void* stream);

/* FP32 peak micro-benchmark used as the roofline denominator of kernel (b) (MEASURED_PEAKS.json has no FP32
 * figure).  Runs `iters` rounds of 8 independent packed FMAs (FFMA2, fma.rn.f32x2) per thread on a full grid and returns
 * the flop count through *flops_host; time it with CUDA events on `stream`.  sink f32 [>=1]. */
int hfrep_ffma_peak(int iters, float* sink, double* flops_host, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HFREP_H */
