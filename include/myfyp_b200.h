/* myfyp_b200 - C ABI of libmyfyp_b200.so (hand-written sm_100a kernels + the native side of the fabric).
 *
 * Conventions: every entry point returns 0 on success and non-zero on failure (b2_last_error() has the text);
 * kernels launch on the CURRENT device into `stream` (a cudaStream_t passed as void*); pointers are device
 * pointers unless the name says host; "slab" = flat fp32 buffer, 16-byte aligned.  Source pointers of the
 * aggregation kernels may be peer-GPU memory (peer access or CUDA-IPC mappings): the loads are the transfer.
 * Each declaration cites the reference (PrivEimantas/myFYP = p2pfl 0.4.0) code whose math it implements.
 */
#ifndef MYFYP_B200_H
#define MYFYP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------------ runtime */
const char* b2_last_error(void);
int b2_version(void);
int b2_device_count(void);
/* cudaDeviceEnablePeerAccess for every ordered GPU pair of this process; returns the number of enabled pairs. */
int b2_enable_peer_access_all(void);
int b2_enable_peer_access(const int* devs, int n); /* peer access among the listed GPUs only; returns the number of enabled ordered pairs */
/* Publish arena of one rank (cudaMalloc + zero fill) and its CUDA-IPC export / import. */
void* b2_arena_alloc(int64_t bytes);
int b2_arena_free(void* p);
int b2_ipc_export(void* base, char* out_handle64);
void* b2_ipc_open(const char* handle64);
int b2_ipc_close(void* p);
/* Interprocess events (cudaEventInterprocess): recorded by the publisher after the copy into an arena slot, waited for by the
 * consumer's stream before the slot is read - the cross-process edge that replaces a host synchronisation.  No reference
 * counterpart (the reference ships pickled bytes, gossiper.py:229-235). */
void* b2_ipc_event_create(void);
int b2_ipc_event_export(void* ev, char* out_handle64);
void* b2_ipc_event_open(const char* handle64);
int b2_event_record(void* ev, void* stream);
int b2_stream_wait_event(void* stream, void* ev);
int b2_event_destroy(void* ev);
/* NVLS probe: per device CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED and, where at least two devices report it, whether a multicast
 * object over them can actually be created (the prerequisite of multimem.ld_reduce / multimem.st aggregation).  Human-readable
 * report in `out`; returns 1 = usable, 0 = not exposed to this process, < 0 = driver error.  No reference counterpart. */
int b2_multicast_probe(char* out, int cap);
int b2_multicast_supported(void); /* attribute of the current device only; creates no context elsewhere */
/* 1 / 0: launch the training-step kernels of the bf16 path with / without programmatic dependent launch (default off; B2_PDL=1). */
void b2_set_pdl(int on);

/* ------------------------------------------------------------------------------------------------ aggregation */
/* A1 FedAvg K-way pull-reduce: out = (accumulate ? out : 0) + sum_k weights[k] * srcs[k].  `srcs`/`weights` are HOST
 * arrays of K <= 64 entries.  Reference math: p2pfl/learning/aggregators/fedavg.py:61-69 (NumPy, per layer). */
int b2_fedavg(float* out, const void* const* srcs, const float* weights, int K, int64_t n, int accumulate, void* stream);
/* A3 streaming accumulate: acc = beta*acc + w*src (the partial-model transfer fused with the accumulation).
 * Reference: aggregator.add_model + fedavg.py:61-69. */
int b2_accumulate(float* acc, const float* src, float w, float beta, int64_t n, void* stream);
/* A4 model pull copy: dst = src (InitModel / FullModel diffusion).  Reference: full_model_command.py:72 + the
 * pickle / deepcopy / load_state_dict chain (lightning_model.py:69-98, p2pfl_model.py:81-101,185). */
int b2_pull_copy(float* dst, const float* src, int64_t n, void* stream);
/* A5 coordinate-wise median (trim < 0; even K -> mean of the two middle values) or trimmed mean (drop the `trim`
 * smallest and largest).  Reference: intended math of the FedMedian stub, fedmedian.py:55-65. */
int b2_median(float* out, const void* const* srcs, int K, int trim, int64_t n, void* stream);
/* A8 Scaffold server update: x += sum_k wy[k]*dy[k];  c += wc * sum_k dc[k].  Reference: scaffold.py:70-111. */
int b2_scaffold_server(float* x, float* c, const void* const* dys, const void* const* dcs, const float* wy, float wc, int K, int64_t n, void* stream);
/* T7 Scaffold client finalize: adj=(x-y)*inv_klr; dc=adj; c_i+=adj; dy=y-x.  Reference: scaffold_callback.py:124-140. */
int b2_scaffold_finalize(float* c_i, float* dy, float* dc, const float* x, const float* y, float inv_klr, int64_t n, void* stream);

/* A2 two-shot FedAvg across G GPUs as one kernel per GPU: reduce-scatter by coordinate shard + all-gather, both as peer
 * loads, synchronised by flags in peer-mapped memory.  Every rank fills one b2_twoshot_args (in DEVICE memory) with the
 * addresses of all ranks' inputs / outputs / flag blocks as mapped on ITS device - once - and launches with the same `epoch`
 * (1, 2, 3... per call).  Flag blocks hold b2_twoshot_flag_bytes() bytes and start zeroed.  `ranks_in_launch` > 1 runs
 * several ranks in one launch on one GPU (dev_args is then an array) - used for single-GPU testing.
 * No reference counterpart (the reference averages pickled NumPy arrays on the host, fedavg.py:61-69). */
#define B2_TWOSHOT_MAX_RANKS 16
typedef struct {
  const void* srcs[B2_TWOSHOT_MAX_RANKS];  /* input slab of every rank */
  void* outs[B2_TWOSHOT_MAX_RANKS];        /* output slab of every rank (outs[rank] is written by this rank) */
  void* flags[B2_TWOSHOT_MAX_RANKS];       /* flag block of every rank */
  int64_t n;                               /* elements, multiple of 4 */
  int world, rank;
} b2_twoshot_args;
int64_t b2_twoshot_flag_bytes(void);
/* `weights` is a HOST array of `world` floats (n_g / sum n); `epoch` must be 1, 2, 3... on successive calls (same on every rank). */
int b2_fedavg_twoshot(const b2_twoshot_args* dev_args, int ranks_in_launch, int64_t n, const float* weights, int world, unsigned int epoch, int max_ctas,
                      void* stream);

/* A2, NVLS variant: the same average through NVSwitch multicast (csrc/nvls.cu): rank r reduces coordinate shard r with
 * multimem.ld_reduce (the switch adds the ranks' values) and multimem.st broadcasts it; barriers are multimem.red on a multicast
 * flag.  Set-up: the root creates the multicast object (b2_nvls_root_create -> POSIX fd, handed to the other ranks), every rank
 * joins (imports + adds its device), and - after ALL ranks have joined - binds its buffer.  Region layout: [flags | in | out],
 * offsets in bytes.  Inputs must be pre-weighted; `scale` multiplies the sum.  Waits are bounded (~8.6 s); b2_nvls_last_timeout
 * reports a call that gave up.  Availability: b2_multicast_probe.  No reference counterpart (fedavg.py:61-69 runs on the host). */
int64_t b2_nvls_granularity(int world);
int b2_nvls_root_create(int world, int64_t bytes, int* out_fd);
void* b2_nvls_join(int fd, int world, int rank, int64_t bytes);
int b2_nvls_bind(void* comm);
void* b2_nvls_local_ptr(void* comm);
void* b2_nvls_mc_ptr(void* comm);
int64_t b2_nvls_flag_bytes(void);
int b2_nvls_allreduce(void* comm, int64_t off_in, int64_t off_out, int64_t n, unsigned int epoch, float scale, int max_ctas, void* stream);
int b2_nvls_last_timeout(void* comm, unsigned int* epoch_out);
int b2_nvls_destroy(void* comm);

/* A6 Krum Gram matrix on tcgen05: gram[i*ld+j] = <src_i - src_0, src_j - src_0>, i,j < K <= 64, ld = 64, fp32 via
 * 3xTF32 split.  Split-K over coordinate chunks: every CTA keeps its partial Gram in TMEM, spills it to
 * `workspace` and the last CTA to finish reduces the partials in a fixed order (deterministic).
 * `workspace` must hold b2_krum_gram_workspace_bytes() bytes and its first 4 bytes must be zero before the first
 * call (the kernel re-zeroes them).  No reference counterpart (Krum is absent from p2pfl). */
int64_t b2_krum_gram_workspace_bytes(void);
int b2_krum_gram(float* gram, const void* const* srcs, int K, int64_t n, void* workspace, int passes, void* stream);
/* Allocation-free variant of b2_krum_gram (no counterpart in the reference): same kernel, same output, but the split-K
 * scratch is a 1 048 832-byte array built into the library image instead of a caller workspace.  For hosts that cannot
 * allocate device memory where they call this (another runtime owns the HBM, or a device-wide synchronising cudaMalloc is
 * not acceptable next to running training streams).  Limits: K <= 64, n <= b2_krum_gram_static_max_n() (2^26); larger
 * requests fail with an error.  The first call on a device resolves the scratch address - issue it before any stream
 * capture.  All calls share the scratch: do not let two of them overlap on the GPU. */
int64_t b2_krum_gram_static_max_n(void);
int b2_krum_gram_static(float* gram, const void* const* srcs, int K, int64_t n, int passes, void* stream);
/* A7 Krum scores + ranking from a K x K squared-distance matrix: scores[i] = sum of the K-f-2 smallest d(i,j), j != i;
 * order[] = indices sorted by (score, index). */
int b2_krum_select(const float* dist, int K, int f, float* scores, int* order, void* stream);

/* ------------------------------------------------------------------------------------------------ local training */
/* T6 fused Adam over a flat slab; g' = g + wd*p + prox_mu*(p - prox_ref) + cv_scale*(cv_global - cv_local); clears g.
 * Reference: torch.optim.Adam (lightning_model.py:181-183), Scaffold term scaffold_callback.py:90-110. */
int b2_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps, float weight_decay, float bias_corr1,
                 float bias_corr2, const float* dev_state, float step_offset, void* p_bf16, float prox_mu, const float* prox_ref, const float* cv_global,
                 const float* cv_local, float cv_scale, int zero_grad, void* stream);
/* Device-resident Adam step state {step, 1-beta1^step, 1-beta2^step}: advance it by one step.  When `dev_state` is passed to
 * b2_adam_step the bias corrections are read from it (so a captured CUDA graph of an epoch replays correctly). */
int b2_adam_prepare(float* state, float beta1, float beta2, void* stream);
/* Graph-friendly stepping: with `dev_state` and step_offset > 0 the step number is dev_state[0] + step_offset (steps finished
 * before this graph replay, kept on the device, plus the index inside the captured epoch, baked into the launch);
 * b2_adam_advance adds the epoch's step count once per replay.  `p_bf16` (optional) receives a bf16 mirror of the updated
 * parameters - the tensor-core operands of the next step; b2_cast_slab (re)builds the whole mirror after a model load. */
int b2_adam_advance(float* state, float n_steps, void* stream);
int b2_cast_slab(const float* p, void* p_bf16, int64_t n, void* stream);
int b2_sgd_step(float* p, float* g, float* buf, int64_t n, float lr, float momentum, float weight_decay, float prox_mu, const float* prox_ref,
                const float* cv_global, const float* cv_local, float cv_scale, int zero_grad, void* stream);
/* T3+T8 fused log-softmax + NLL (+ gradient, + argmax accuracy).  loss_out/correct are accumulated atomically.
 * Reference: F.cross_entropy(log_softmax(x)) lightning_model.py:178,189 and torchmetrics Accuracy :204. */
int b2_xent_fwd_bwd(const float* logits, const int64_t* labels, float* loss_out, float* dlogits, float* correct, int B, int C, float scale, void* stream);

/* Training variant in one block: loss written, dlogits as bf16 (pitch ld16) for the backward GEMMs, bias gradient added. */
int b2_xent_train(const float* logits, const int64_t* labels, float* loss_out, void* dlogits_bf16, int ld16, float* bias_grad, int B, int C, float scale, void* stream);
/* confusion[label*C + argmax(logits)] += 1 (fp32 counts) - statistic behind accuracy / precision / recall / F1.
 * Reference: the fork author's extended MLP metrics (mlp_pytorch.txt, torchmetrics F1/Precision/Recall). */
int b2_confusion(const float* logits, const int64_t* labels, float* conf, int B, int C, void* stream);

/* T1/T2 bf16 GEMM on tcgen05 + TMA: D[M,N] = epilogue(A[M,K] * B[N,K]^T).  See csrc/gemm.cu for the epilogue flags.
 * Reference: the three nn.Linear GEMMs of the MLP and their autograd backward (lightning_model.py:169-179). */
typedef struct {
  const void* A;      /* bf16 [M, lda] (K-major) or, with a_mn, [K, lda] */
  const void* B;      /* bf16 [N, ldb] (K-major) or, with b_mn, [K, ldb] */
  int M, N, K;        /* any sizes (TMA zero-fills the tails); lda/ldb in elements, multiples of 8 */
  int lda, ldb;
  float* out_f32;     /* optional fp32 [M, ldo] */
  void* out_bf16;     /* optional bf16 [M, ldo16] */
  void* out_bf16_t;   /* optional bf16 transposed [N, ldt] (activation kept K-major for the next wgrad) */
  int ldo, ldo16, ldt;
  const float* bias;  /* optional per-N bias (added before the activation) */
  const void* mask;   /* optional bf16 [M, ldm]: out *= (mask > 0)  (ReLU backward) */
  int ldm;
  float* colsum;      /* optional fp32 [N]: += column sums of the output (bias gradient), atomic */
  int relu;           /* apply ReLU */
  int accumulate;     /* out_f32 += result instead of = */
  int a_mn, b_mn;     /* operand stored transposed: A as [K, lda>=M], B as [K, ldb>=N] (MN-major UMMA operands) */
  int split_k;        /* > 1: split the K loop over grid.z, atomically adding into out_f32 (which must hold 0 or the running sum) */
  float* colsumsq;    /* optional fp32 [N]: += column sums of squares (batch-norm statistics of the output), atomic */
  int perm_cin, perm_taps; /* > 0: out_f32 column j = tap * cin + c is stored at c * taps + tap - the weight gradient of a convolution
                              whose patch matrix is in (tap, c) order lands in PyTorch's [Cout, Cin, KH, KW] layout */
} b2_gemm_args;
int b2_gemm_bf16(const b2_gemm_args* args, void* stream);

/* Implicit-GEMM convolution on the same tcgen05 kernel (stride 1, pad = ksize / 2, cin % 64 == 0): the patch matrix is never
 * written - 4-D TMA boxes of the NHWC activation, shifted per filter tap and zero-filled outside the image, are the operand tiles.
 * `weights` is bf16 [cout, ksize*ksize*cin] in (kh, kw, cin) order (b2_repack_conv_weights).  Epilogue fields as in b2_gemm_args.
 *   B2_CONV_FWD    act = x  [batch,h,w,cin]   -> out [batch*h*w, cout]   (+ colsum / colsumsq: batch-norm statistics)
 *   B2_CONV_DGRAD  act = dz [batch,h,w,cout]  -> out [batch*h*w, cin]    (+ add_bf16: gradient arriving through a second consumer)
 *   B2_CONV_WGRAD  act = x, dz = [batch*h*w, cout] -> out_f32 [cout, cin, ksize, ksize] (PyTorch's layout), split-K over pixels
 * Reference path: torch conv2d through autograd / cuDNN (the reference ships no CNN; BASELINE.json configs 2-4 name CNN / ResNet-18). */
#define B2_CONV_FWD 0
#define B2_CONV_DGRAD 1
#define B2_CONV_WGRAD 2
typedef struct {
  int mode;
  const void* act;      /* bf16 NHWC 4-D operand */
  const void* weights;  /* bf16 [cout, taps*cin] (fwd, dgrad) */
  const void* dz;       /* bf16 [batch*h*w, cout] (wgrad) */
  int batch, h, w, cin, cout, ksize;
  float* out_f32;
  void* out_bf16;
  const float* bias;
  const void* add_bf16;
  float* colsum;
  float* colsumsq;
  int relu, accumulate, split_k;
} b2_conv_gemm_args;
int b2_conv_gemm(const b2_conv_gemm_args* args, void* stream);
/* fp32 [cout, cin, taps] (PyTorch conv weight inside the slab) -> bf16 [cout, taps, cin] for every layer of `table` in one launch;
 * table = n_layers x {src offset, dst offset, cout, cin, taps, dst row pitch >= taps*cin} (int64, device memory). */
int b2_repack_conv_weights(const float* slab, void* dst_bf16, const int64_t* table, int n_layers, int64_t max_elems, void* stream);
/* fp32 [R,C] (pitch lds) -> bf16 row-major copy (pitch ldd, optional), bf16 transposed copy [C,R] (pitch ldt, optional),
 * column sums += (optional, atomic).  Staging step of the GEMMs above; the reference has no counterpart (it trains in fp32
 * through torch autograd). */
int b2_cast_transpose(const float* src, int R, int C, int lds, void* dst_bf16, int ldd, void* dst_bf16_t, int ldt, float* colsum, void* stream);

/* T4/T5 convolution building blocks (NHWC bf16 activations = the [pixels, channels] matrices b2_gemm_bf16 reads and writes).
 * Patch index order is (kh, kw, c): a tap's channels are contiguous like the activations (16-byte vector gathers); weights are
 * repacked to bf16 [Cout, taps, Cin] by b2_repack_conv_weights and the weight-gradient GEMM scatters back (perm_cin / perm_taps).
 * Reference path: torch conv2d / max_pool2d via autograd (the reference ships no CNN; BASELINE.json configs 2-4 name one). */
int b2_im2col(const void* x, int x_is_nchw_f32, void* col_bf16, int ldc, int B, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream);
/* dx = col2im(dcol) (+ add, when the activation has a second consumer such as a residual / shortcut branch) */
int b2_col2im(const void* dcol_bf16, int ldc, void* dx_bf16, const void* add_bf16, int B, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream);
/* T5 batch normalisation on [M, C] NHWC bf16.  Training: `stats` = [sum[C], sumsq[C]] as accumulated by b2_gemm_bf16 (colsum / colsumsq);
 * y = act(gamma (z - mean) invstd + beta (+ residual)); running statistics and the batch counter are updated with PyTorch semantics.
 * Eval: stats == NULL, the running statistics are used.  Reference path: torch batch_norm via autograd (ResNet-18 is not in the reference). */
int b2_bn_apply(const void* z_bf16, const void* residual_bf16, void* y_bf16, const float* stats, const float* gamma, const float* beta, float* running_mean, float* running_var,
                float* num_batches, float eps, float momentum, int M, int C, int relu, void* stream);
/* g = dy * (y > 0 if relu) (stored: it is also the gradient of the residual branch); dz = gamma invstd (g - mean g - xhat mean(g xhat));
 * dgamma += sum g xhat, dbeta += sum g.  `red2c` is a 2C-float scratch (zeroed here unless the caller says it already is:
 * a training step zeroes one arena for all its layers).  C % 8 == 0, C <= 2048. */
int b2_bn_backward(const void* dy_bf16, const void* y_bf16, const void* z_bf16, void* g_bf16, void* dz_bf16, float* red2c, const float* stats, const float* gamma, float* dgamma,
                   float* dbeta, float eps, int M, int C, int relu, int red_is_zero, void* stream);
int b2_avgpool_fwd(const void* x_bf16, void* out_bf16, int ldo, int B, int HW, int C, void* stream);
int b2_avgpool_bwd(const void* dout_bf16, int ldo, void* dx_bf16, int B, int HW, int C, void* stream);
int b2_maxpool2_fwd(const void* y_bf16, void* p_bf16, void* idx_u8, int B, int H, int W, int C, void* stream);
/* dz = dp routed to the remembered maximum, times the ReLU mask (y > 0) of the pre-pool activation */
int b2_maxpool2_relu_bwd(const void* dp_bf16, const void* idx_u8, const void* y_bf16, void* dz_bf16, int B, int H, int W, int C, void* stream);
int b2_colsum_bf16(const void* z_bf16, int M, int C, int ld, float* out, void* stream);
/* Fused forward tail: 2x2 max-pool of y [B,2PH,2PW,C] (winner index kept) scattered straight into the NEXT layer's col matrix
 * (geometry KH,KW,stride,pad over the pooled [B,PH,PW,C] map); the pooled tensor is never materialised. */
int b2_pool_im2col(const void* y_bf16, void* idx_u8, void* col_bf16, int ldc, int B, int PH, int PW, int C, int KH, int KW, int stride, int pad, void* stream);
/* Fused backward head: col2im of the next layer's dcol -> unpool to the remembered winner -> ReLU mask -> dz [B,2PH,2PW,C],
 * bias_grad[c] += column sums of dz. */
int b2_unpool_col2im(const void* dcol_bf16, int ldc, const void* idx_u8, const void* y_bf16, void* dz_bf16, float* bias_grad, int B, int PH, int PW, int C, int KH, int KW,
                     int stride, int pad, void* stream);

/* ------------------------------------------------------------------------------------------------ node-batched TF32 training
 * One training step of a conv3x3-ReLU-maxpool2 ... Flatten-Linear-ReLU-Linear network for G co-resident federated nodes per
 * launch sequence (csrc/train_tf32.cu): fp32 storage, convolutions on tcgen05 kind::tf32 (torch's default precision for
 * cuDNN convolutions), linear layers as 3xTF32 (fp32-class accuracy, torch's default for nn.Linear), loss / head in fp32.
 * Every kernel takes the node ("slot") index from blockIdx.y; tensor maps, shapes and epilogue pointers of each slot sit in
 * device memory (built once by b2_cnn_plan_create), per-node pointers in a bind table rewritten before a graph replay.
 * Replaces the reference's one-actor-process-per-learner pool (p2pfl/learning/frameworks/simulation/actor_pool.py:101-128,
 * virtual_learner.py:101-137) and its Lightning training loop (pytorch/lightning_learner.py:78-100). */
#define B2_CNN_MAX_CONV 4
typedef struct {
  int groups;   /* G: slots (nodes trained together) */
  int batch;    /* samples per step and node */
  int image;    /* input height = width */
  int in_ch;    /* 1 or 3 */
  int n_conv;   /* 2..4 conv blocks */
  int conv_ch[B2_CNN_MAX_CONV];
  int hidden, classes;
  int64_t off_conv_w[B2_CNN_MAX_CONV], off_conv_b[B2_CNN_MAX_CONV]; /* element offsets inside the parameter slab */
  int64_t off_fc1_w, off_fc1_b, off_fc2_w, off_fc2_b;
  int64_t slab_elems; /* trainable elements of the slab (multiple of 4) */
} b2_cnn_desc;
typedef struct {
  float* slab;          /* the node's fp32 parameters */
  float* m;             /* Adam moments */
  float* v;
  float* adam_state;    /* {steps done, ...} as b2_adam_step's dev_state */
  const float* x;       /* [n, in_ch, image, image] fp32 */
  const int64_t* y;     /* [n] */
  const int64_t* perm;  /* [n] sample order, or NULL = identity */
  float* losses;        /* training: [steps] mean loss per step (+=); evaluation: {sum nll, #correct} (+=) */
  const float* prox_ref;   /* FedProx anchor or NULL */
  const float* cv_global;  /* Scaffold c or NULL */
  const float* cv_local;   /* Scaffold c_i */
  float prox_mu, cv_scale, lr, weight_decay;
  int n_samples;
  int active;
  int sgd;          /* 0: Adam (m, v);  1: SGD with momentum (buffer in m, v unused) */
  float momentum;
} b2_cnn_bind;
int64_t b2_cnn_workspace_bytes(const b2_cnn_desc* d);
/* `workspace` (device, 1024-byte aligned, b2_cnn_workspace_bytes) is carved into every slot's activations, weight mirrors and
 * gradient mirror plus the problem / bind tables; the call encodes all tensor maps and uploads the tables on `stream`. */
void* b2_cnn_plan_create(const b2_cnn_desc* d, void* workspace, int64_t bytes, void* stream);
int b2_cnn_plan_destroy(void* plan);
int b2_cnn_plan_bind(void* plan, const b2_cnn_bind* host_binds, void* stream);   /* G entries; async copy through pinned staging */
int b2_cnn_plan_refresh(void* plan, void* stream);                                /* slab -> tf32 weight mirrors, clear gradient mirrors */
int b2_cnn_plan_train_step(void* plan, int step, void* stream);                   /* all kernels of one step for all active slots; returns <0 on error else #launches */
int b2_cnn_plan_eval_step(void* plan, int step, void* stream);                    /* forward + loss / accuracy accumulation */
int b2_cnn_plan_advance(void* plan, int steps, void* stream);                     /* adam_state[0] += steps the slot really took (once per epoch) */
/* test / debug: run one kernel of the step (`which` as enumerated in train_tf32.cu), look a slot buffer up */
int b2_cnn_plan_run(void* plan, int which, int step, void* stream);
int b2_cnn_plan_query(void* plan, int slot, int buf, void** ptr, int64_t* bytes);

#ifdef __cplusplus
}
#endif
#endif /* MYFYP_B200_H */
