/* rgp.h — C-ABI of librgp_sm100.so: the B200 (sm_100a) hot path of Recurrent Gaze Prediction.
 *
 * The reference (yj-yu/Recurrent_Gaze_Prediction) has no FFI layer: every numeric site is a
 * TensorFlow-1 / NumPy library call made from Python.  Each entry point below names the
 * reference call site(s) it replaces (paths relative to the reference root).  All pointers are
 * DEVICE pointers owned by the caller unless a comment says "host"; `stream` is a cudaStream_t
 * passed as void*; every function returns 0 on success or a negative code, never throws, never
 * allocates device memory (workspace sizes are queried / documented per op) and records a
 * message retrievable with rgp_last_error().  dtype: 0 = fp32 operands (CUDA-core FMA engine),
 * 1 = bf16 operands with fp32 accumulation (tcgen05 / TMEM / TMA engine).
 *
 * Row layouts ("q-layout" / "p-layout"), see DESIGN.md section 3:
 *   a G x G frame is stored as Wp = G+1 columns and G+1 rows of C-channel vectors (frame stride
 *   S = Wp*Wp rows); position (y,x) lives at p-row  f*S + (y+1)*Wp + (x+1)  (operand layout,
 *   zero everywhere else) and produces its output at q-row f*S + y*Wp + x.  A 3x3 SAME conv tap
 *   (ky,kx) is then the constant row shift ky*Wp + kx, i.e. a plain 2-D TMA box.
 */
#ifndef RGP_H_
#define RGP_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define RGP_MAX_TAPS 32   /* 25 = the 5x5 top cell of the cascade (gaze_grcn_cascade.py:346-351) */

/* epilogue selectors of rgp_tapgemm */
enum {
  RGP_EPI_STORE_F32 = 0,   /* out_f32[m*ldo + n] (+)= acc                                          */
  RGP_EPI_PROJ = 1,        /* dense row -> p-row, + bias, store operand dtype (gaze_grcn.py:239-242) */
  RGP_EPI_GRU_ZR = 2,      /* z = sig(acc+Xz) ; r = sig(acc+Xr) ; rh = r*h   (gaze_grcn.py:108-119)  */
  RGP_EPI_GRU_C = 3,       /* c = tanh(acc+Xc); h' = z*h + (1-z)*c ; BN-scale (gaze_grcn.py:122-127,325) */
  RGP_EPI_HEAD = 4,        /* composed transposed-conv head scatter + out_b  (gaze_grcn.py:326-366)  */
  RGP_EPI_STORE_OP_P = 5,  /* store operand dtype at p-row (valid rows only): dX of a SAME conv       */
  RGP_EPI_STORE_OP = 6,    /* store operand dtype at row m                                            */
  /* BPTT (base.py:278-281 tf.gradients of gaze_grcn.py:108-127) */
  RGP_EPI_BWD_RH = 7,      /* acc = d(r.h): dpre_r = acc*h*r(1-r) -> dG[:,Ch:2Ch]; dhpart += acc*r     */
  RGP_EPI_BWD_H = 8,       /* acc = convT(dpre_z,dpre_r): dh = acc + dhpart + head term; gate backward of step t-1 */
  RGP_EPI_DHS = 9,         /* head backward: store fp32 acc at the recurrence q-row of the head-layout row */
  RGP_EPI_DX_DENSE = 10,   /* store operand dtype at the dense row f*P + y*G + x (valid rows only)     */
  /* cascade (gaze_grcn_cascade.py:317-336): the 11x11 / stride-7 SAME transposed conv 7x7 -> 49x49 as a 4-tap GEMM over the
   * 8x8 p-layout rows of h_t.  Row (a,b) of frame f and column n = (ry*7+rx)*P + co is output pixel
   * (oy,ox) = (7a+ry-2, 7b+rx-2), channel co; it is stored (operand dtype) at row f*So + (oy+2)*Wo + (ox+2) of out_op
   * (leading dimension ldo) - the zero-bordered operand layout of the 5x5 top cell; columns n >= 49*P are padding. */
  RGP_EPI_UP7 = 11,
  /* the cascade's up-sampling AND the input half of its 5x5 top cell as ONE 3x3-tap GEMM over the 7x7 grid (there is no
   * nonlinearity between gaze_grcn_cascade.py:326-331 and the cell's input convs, rnn_cell GRU_RCN_Cell fed at :379):
   * position-major tiles only (im2col = 3: M = frames*49, A = the bottom states, p-layout S/Wp/G = 7); Bw = the composed
   * weights of rgp_cascade_compose, [class(y)*3 + class(x)][tap][N = 1024][K], class = 0 / 1 / 2 for the first / inner /
   * last block row or column; column n = (py*7 + px)*16 + gate of block (y, x) is stored at row
   * f*So + (7y+py)*Wo + 7x+px, column gate of out_f32 (leading dimension ldo >= 16) - the xg of rgp_topgru_seq_fwd. */
  RGP_EPI_UP7X = 12
};

/* One "tap GEMM":  acc[m,n] = sum_tap sum_k A[m + tap_off[tap], k] * Bw[tap*N + n, k]
 * A: [rowsA, lda] operand dtype, row-major.  Bw: [ntaps*N, K] operand dtype (K contiguous).
 * This single contraction serves the 1x1 projection, the 3x3 SAME convolutions of the ConvGRU /
 * ConvLSTM cells (input half for all T frames at once, recurrent half per step) and the
 * composed transposed-convolution head. */
typedef struct rgp_tapgemm_args {
  const void* A; long long rowsA; int lda;
  const void* Bw;
  int M, N, K, ntaps;
  int tap_off[RGP_MAX_TAPS];
  int dtype;            /* 0 fp32, 1 bf16 */
  int engine;           /* 0 = by dtype (fp32 -> FMA engine, bf16 -> tcgen05 engine); 1 = force the FMA engine */
  int epi;              /* RGP_EPI_* */
  int accumulate;       /* STORE_F32 only: out += acc */
  /* geometry of the row layouts */
  int S, Wp, G;         /* q-layout of the M rows: frame stride, width, valid extent */
  int P;                /* PROJ: dense rows per frame (G*G) */
  int So, Wo, pado;     /* secondary output layout (PROJ: p-layout S,Wp,1; GRU_C: head input layout) */
  int Ch;               /* hidden width */
  int B, T, t;          /* clips, frames per clip, current step (GRU_C / HEAD time-major -> [B,T]) */
  /* generic pointer slots; meaning per epilogue is documented in DESIGN.md section 4 */
  float* out_f32; int ldo;
  void* out_op;  void* out_op2;
  const float* xg; int ldx;     /* precomputed input-half pre-activations, q-layout */
  const float* hprev_f32;       /* p-layout fp32 master state */
  float* hnext_f32;
  float* z_f32;                 /* q-layout [M, Ch] */
  float* save_r; float* save_c; /* optional (training): q-layout [M, Ch] */
  const float* bias;            /* PROJ: [N]; HEAD: [1]; GRU_C: bn scale [Ch] (may be NULL) */
  const float* bias2;           /* GRU_C: bn shift [Ch] (may be NULL) */
  float alpha;                  /* GRU_C: 1/sqrt(1+eps) of the frozen batch-norm (gaze_grcn.py:325) */
  /* BPTT slots */
  float* dhpart;                /* [M, Ch] fp32 q-layout: part of dL/dh_{t-1} that does not pass through a conv */
  const float* dhs;             /* [M, Ch] fp32 q-layout: dL/d(head input) of the step whose gates are differentiated */
  const float* save_z;          /* [M, Ch] saved z of that step */
  void* dg; int ldg;            /* dL/d(pre-activations) [M, 3Ch] operand dtype, columns z | r | c */
  /* Dense-M mode of a 3x3 SAME convolution (RGP_EPI_STORE_F32, 9 taps in (ky,kx) order, tcgen05 engine only): M counts only
   * the G*G real positions of each frame (M = frames * G*G instead of frames * S: 49 instead of 64 rows per 7x7 frame).  A
   * is the same p-layout operand (S, Wp, G describe it); each tap's A tile is fetched by a TMA im2col load with hardware
   * zero fill at the frame borders, and row m = f*G*G + y*G + x is stored at the q-row f*S + y*Wp + x of out_f32.
   * im2col = 2: the TRANSPOSE of that convolution (BPTT: dX of the input half) with RGP_EPI_DX_DENSE - A is the q-layout
   * gradient (zero at the pad positions), tap (ky,kx) reads position (y+1-ky, x+1-kx), row m is stored at dense row m.
   * im2col = 3 / 4: the same two convolutions with POSITION-MAJOR tiles (one tile = one position (y,x) of 128 consecutive
   * frames, so out-of-frame taps are skipped for the whole tile: (3G-2)^2 of 9 G^2 tap-tiles); same arguments, same
   * results bit for bit.  The library takes this order by itself for im2col = 1 / 2 when it is cheaper (many frames);
   * RGP_PMAJOR=0 in the environment keeps the dense-M order. */
  int im2col;
} rgp_tapgemm_args;

int rgp_tapgemm(const rgp_tapgemm_args* a, void* stream);

/* Persistent ConvGRU recurrence (gaze_grcn.py:95-129 unrolled by :259-288): ALL T steps of the hidden half
 * in ONE launch.  A CTA owns 128 consecutive q-rows (two clips at G=7: frames never straddle a tile, so clips
 * in different tiles never exchange data) and walks t = 0..T-1 by itself: per step three tcgen05 tap-GEMMs
 * (U_r*h, U_z*h, U*(r.h)) with the gate maths fused between them; the state never leaves the SM's reach
 * (fp32 master + bf16 operand rows in L2), there is no per-step launch and no grid-wide barrier.
 * bf16 operands only (dtype 1).  Requires 128 % S == 0 (G = 7, 3 or 1) and Ch in {128, 256}.
 * Buffers are those of the per-step path (rgp_tapgemm RGP_EPI_GRU_ZR / RGP_EPI_GRU_C), see DESIGN.md section 3. */
typedef struct rgp_gru_seq_args {
  void* Hop; long long rowsH;     /* bf16 p-layout state, slab t = h_t (t = 0..T), slab stride B*S rows        */
  float* Hf;                      /* fp32 master of the same                                                     */
  void* RH; long long rowsRH;     /* bf16 p-layout r.h; one slab (inference) or T slabs (train)                  */
  const void* Uzr;                /* [9*2Ch, Ch] bf16 op layout (tap, z|r out channel, in channel)               */
  const void* Uc;                 /* [9*Ch, Ch]                                                                  */
  const float* Xg;                /* [T*B*S, 3Ch] fp32 input-half pre-activations (columns z | r | c)            */
  float* Zs; float* Rs; float* Cs;/* train: saved gates [T, B*S, Ch] fp32 q-layout each; inference: all NULL       */
  void* HS;                       /* bf16 head-input layout [T*B*So, Ch] (BN-scaled h_t), or NULL                */
  const float* bn_gamma; const float* bn_beta; float alpha;   /* [T, Ch] each, or NULL (gaze_grcn.py:325)       */
  int B, T, G, Ch;
  int So, Wo, pado;               /* head-input layout geometry                                                  */
  int nsplit;                     /* 1, or 2: a 2-CTA cluster shares a tile, each CTA owns half of the channels  */
  int dtype;                      /* must be 1 (bf16 operands, tcgen05 engine)                                   */
  int mcast;                      /* 1, 2, 4 or 8 tiles per cluster: each CTA fetches 1/mcast of every weight box
                                     and TMA-multicasts it to the others (nsplit*mcast <= 8)                      */
  int ntile;                      /* 0 / 1: one 128-row tile per CTA; 2 (needs Ch/nsplit = 128, mcast = 1): a CTA walks two
                                     tiles for its channel half, every streamed weight box feeds both (half the L2->SMEM
                                     bytes per MMA)                                                                 */
} rgp_gru_seq_args;
int rgp_convgru_seq_fwd(const rgp_gru_seq_args* a, void* stream);
/* Profiling hook: phase timestamps (globaltimer ns) CTA 0 recorded during the last rgp_convgru_seq_fwd, copied to a
 * HOST array of n <= 32*12 values laid out [step][event] (events listed in csrc/gru_seq.cu).  Synchronises the device. */
int rgp_convgru_seq_profile(unsigned long long* host_out, int n);

/* Small-batch persistent ConvGRU (same cell, gaze_grcn.py:95-129 / :259-288; csrc/lstm_seq.cu): CTA (tile, split) owns 128
 * q-rows x 16 channels for all T steps, its slices of U_z | U_r and U (110 KB at hidden 128, the reference's own width) stay in
 * shared memory, h and z of its rows in registers; the two halo exchanges of a step (r.h, then h_t) go through per-tile
 * counters (no grid barrier, cooperative launch).  For the batches rgp_convgru_seq_fwd is too coarse for (tiles * Ch/16 <=
 * SMs: up to 36 clips at 7x7; the reference trains at 28).  Same argument struct (nsplit / mcast / ntile ignored); `flags`:
 * workspace of 2 * mtiles + 1 int32, ZEROED by the caller before every launch (flags[mtiles] != 0 afterwards = time-out). */
int rgp_convgru_seq_small_fwd(const rgp_gru_seq_args* a, int* flags, void* stream);
int rgp_convgru_seq_small_fits(int B, int G, int Ch);
/* Workspace queries (no hidden allocation - the library never calls cudaMalloc).  Bytes of the int32 counter
 * workspace `flags` the persistent small-batch kernels need (zeroed by the caller before every launch); every other
 * entry point works in place on caller-owned buffers and needs no workspace (0). */
long long rgp_convgru_seq_small_workspace_bytes(int B, int G);   /* forward and BPTT: 2 * mtiles + 1 counters */
long long rgp_convlstm_seq_workspace_bytes(int B, int G);        /* mtiles + 1 counters */
/* ... and its reverse-time mirror: BPTT through the T ConvGRU steps (tf.gradients of gaze_grcn.py:108-127, base.py:278-281)
 * in ONE launch - what the per-step path does with rgp_gru_bwd_last + 2 rgp_tapgemm (RGP_EPI_BWD_RH / RGP_EPI_BWD_H) per step.
 * Fills dG [T*M, 3Ch] (operand dtype, q-layout, columns dpre_z | dpre_r | dpre_c) for the weight-gradient contractions. */
typedef struct rgp_gru_bwd_args {
  void* dG; long long rowsG;      /* [T*B*S, 3Ch] bf16, zero at the pad positions (never written)                          */
  const void* UcT;                /* [9*Ch, Ch]  bf16: rgp_transpose_taps of Uc  (row tap*Ch + in channel, K = out channel) */
  const void* UzrT;               /* [9*Ch, 2Ch] bf16: rgp_transpose_taps of Uzr (K = z | r out channels)                   */
  const float* Zs; const float* Rs; const float* Cs;   /* saved gates [T, B*S, Ch] fp32 q-layout                          */
  const float* Hf;                /* fp32 p-layout states, slab t = input state of step t                                  */
  const float* dhs;               /* [T*B*S, Ch] fp32 q-layout: dL/d(head input) of every step                             */
  const float* bn_gamma; float alpha;   /* [T, Ch] or NULL (gaze_grcn.py:325)                                             */
  int B, T, G, Ch;
  int dtype;                      /* must be 1                                                                             */
} rgp_gru_bwd_args;
/* flags: 2 * mtiles + 1 int32, zeroed by the caller before every launch (flags[mtiles] != 0 afterwards = time-out) */
int rgp_convgru_seq_small_bwd(const rgp_gru_bwd_args* a, int* flags, void* stream);

/* Persistent ConvLSTM recurrence (gaze_lstm.py:103-133 as written, unrolled by :260-286): ALL T steps in ONE launch, for
 * any grid (frames may straddle the 128-row tiles: 14x14, BASELINE config 4).  CTA (tile, split) owns 128 q-rows and
 * Ch/nsplit = 16 channels of all three hidden convolutions for the whole sequence: its slice of W_h (9 taps x 3 gates x
 * 16 output channels x Ch, 110 KB at Ch = 128) is loaded into shared memory ONCE and stays there; per step it fetches
 * the 128 + 2*Wp+2 state rows it needs as one TMA slab, runs 9 x Ch/16 tcgen05 MMAs (N = 48) into TMEM and applies the
 * gate maths in the epilogue (cell state in registers, peepholes in shared memory).  A step's halo rows come from the
 * two neighbouring tiles: tiles publish "step t written" through per-tile counters in `flags` (release/acquire at gpu
 * scope + async-proxy fences), so there is no grid-wide barrier and no per-step launch.  All CTAs must be co-resident
 * (cooperative launch): mtiles * nsplit <= SMs.  bf16 operands (dtype 1) only.
 * Buffers are those of the per-step path (rgp_tapgemm STORE_F32 + rgp_lstm_gates). */
typedef struct rgp_lstm_seq_args {
  void* Hop; long long rowsH;     /* bf16 p-layout state, slab t = h_t (t = 0..T), slab stride B*S rows                  */
  float* Hf;                      /* fp32 master of the same                                                               */
  float* Cf;                      /* fp32 cell, p-layout [B*S + tail, Ch]: in = c_0, out = c_T                             */
  const void* Wh;                 /* [9*3Ch, Ch] bf16 op layout: row tap*3Ch + gate*Ch + out channel (gates i | f | o)     */
  const float* Xg;                /* [T*B*S, 4Ch] fp32 input-half pre-activations, columns i | f | c | o                   */
  const float* wci; const float* wcf; const float* wco;   /* peepholes [G, G, Ch] fp32                                    */
  void* HS;                       /* bf16 head-input layout [T*B*So, Ch] or NULL                                           */
  float* save;                    /* train: [T, B*S, 5Ch] fp32 (i, f, g, o, c_{t-1}) as rgp_lstm_gates stores it, or NULL   */
  int* flags;                     /* workspace: mtiles + 1 int32, ZEROED by the caller before every launch; after the launch
                                     flags[mtiles] != 0 reports a dependency time-out (results invalid)                  */
  int B, T, G, Ch;
  int So, Wo, pado;
  int dtype;                      /* must be 1                                                                             */
} rgp_lstm_seq_args;
int rgp_convlstm_seq_fwd(const rgp_lstm_seq_args* a, void* stream);
/* 1 if rgp_convlstm_seq_fwd can run this shape on the current device (all mtiles * Ch/16 CTAs co-resident), else 0 */
int rgp_convlstm_seq_fits(int B, int G, int Ch);
/* Profiling hook: phase timestamps (globaltimer ns) CTA 0 recorded during the last rgp_convlstm_seq_fwd, copied to a
 * HOST array of n <= 32*8 values laid out [step][event] (events listed in csrc/lstm_seq.cu).  Synchronises the device. */
int rgp_convlstm_seq_profile(unsigned long long* host_out, int n);

/* BPTT mirror of rgp_convlstm_seq_fwd for the co-resident regime (tiles * Ch/16 <= SMs; rgp_convgru_seq_small_fits): the gate
 * backward of every step (rgp_lstm_gates_bwd's arithmetic, gaze_lstm.py:112-133 as written) and the transposed hidden
 * convolution dh_{t-1} = conv^T(dHc_t; W_hi | W_hf | W_ho) for all T steps in ONE launch; fills dG4 and dHc for the weight-
 * gradient contractions and ACCUMULATES the peephole gradients (once per row, at the end). */
typedef struct rgp_lstm_bwd_args {
  void* dHc; long long rowsHc;    /* [T*B*S, 3Ch] bf16 q-layout: d(hidden conv outputs) (i+c) | f | o                     */
  void* dG4;                      /* [T*B*S, 4Ch] bf16 q-layout: d(pre-activations) i | f | c | o                         */
  const void* WhT;                /* [9*Ch, 3Ch] bf16: rgp_transpose_taps of Wh (row tap*Ch + in channel, K = 3Ch)        */
  const float* save;              /* [T, B*S, 5Ch] fp32: (i, f, g, o, c_{t-1}) stored by the forward pass                 */
  const float* Cf;                /* fp32 p-layout cell after the last step (c_T)                                         */
  const float* dhs;               /* [T*B*S, Ch] fp32 q-layout: dL/d(head input)                                          */
  const float* wci; const float* wcf; const float* wco;   /* peepholes [G, G, Ch]                                         */
  float* gwci; float* gwcf; float* gwco;                  /* their gradients, accumulated                                 */
  int B, T, G, Ch;
  int dtype;                      /* must be 1                                                                            */
} rgp_lstm_bwd_args;
/* flags: rgp_convlstm_seq_workspace_bytes(B, G) of int32, zeroed by the caller before every launch */
int rgp_convlstm_seq_small_bwd(const rgp_lstm_bwd_args* a, int* flags, void* stream);

/* ---- model path ------------------------------------------------------------------------- */
/* gaze_grcn.py:225-227 (tf.transpose NCHW->NHWC of the C3D blob) fused with the fp32->operand
 * cast and the [B,T]->time-major reorder: c3d [B,T,C,P] fp32 -> out [(t*B+b)*P + p, C]. */
int rgp_pack_c3d(const float* c3d, void* out, int B, int T, int C, int P, int dtype, void* stream);
/* the same with the fed features in src_dtype (0 fp32 as the reference feeds them, gaze_rnn.py:520-531; 1 bf16: half the
 * host->device bytes — the bf16 engine rounds the features to bf16 in this pass anyway, so `out` is bit-identical) */
int rgp_pack_c3d_src(const void* c3d, int src_dtype, void* out, int B, int T, int C, int P, int dtype, void* stream);
/* embedded features [B,T,G,G,C] fp32 (BASELINE 'L1' entry, after the projection) -> p-layout */
int rgp_pack_emb(const float* emb, void* out, int B, int T, int G, int C, int dtype, void* stream);
/* fp32 -> operand dtype cast of a flat buffer (weights after an optimiser step) */
int rgp_cast(const float* src, void* dst, long long n, int dtype, void* stream);
/* gaze_grcn77.py:206-211: logits[b,t,y,x] = h . out_W + out_b from the p-layout fp32 state */
int rgp_fc_head(const float* h_f32, const float* out_W, const float* out_b, float* logits,
                int B, int T, int t, int G, int Ch, void* stream);
/* the same for all T steps in one launch: h1_f32 = slab of h_1 (slab t = h_{t+1}); Ch % 4 == 0, 16-byte aligned */
int rgp_fc_head_all(const float* h1_f32, const float* out_W, const float* out_b, float* logits, int B, int T, int G,
                    int Ch, void* stream);
/* ConvLSTM gate maths, gaze_lstm.py:112-133 (bugs preserved), after the hidden conv GEMM */
int rgp_lstm_gates(const float* xg, int ldx, const float* hconv, const float* wci, const float* wcf,
                   const float* wco, float* c_f32, float* h_f32, void* h_op_next, void* hs_op,
                   int B, int G, int Ch, int So, int Wo, int pado, int dtype, float* save, void* stream);
/* head weight composition (weight space, once per weight version): gaze_grcn.py:292-314,326-366.
 * w1 [5,5,64,Ch], w2 [5,5,32,64], w3 [7,7,12,32], out_w [12] fp32 (TF layouts) ->
 * w123 [19,19,Ch] fp32 (workspace w23 [11,11,64], w3f [7,7,32]) and the operand-dtype tap matrix
 * bhead [16*48, Ch]. */
int rgp_head_compose(const float* w1, const float* w2, const float* w3, const float* out_w,
                     float* w3f, float* w23, float* w123, void* bhead, int Ch, int dtype, void* stream);
/* rgp_head_compose for callers that cannot allocate device memory (CUDA-graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same
 * composition, same bhead [16*48, Ch] (operand dtype) bit for bit; the intermediates w3f, w23 and w123 live in a
 * file-scope device array of the library (776 960 bytes, held by every GPU that loads the library until exit) instead of
 * caller workspaces, so the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 * Limit: Ch <= 512 - larger Ch returns -22 with a message in rgp_last_error().  The array is shared by every call on a
 * device and keeps the last call's contents (each call overwrites all it reads, nothing needs clearing); calls that
 * overlap on the GPU (different streams, no ordering) race on it - order them on one stream or with events. */
int rgp_head_compose_static(const float* w1, const float* w2, const float* w3, const float* out_w, void* bhead, int Ch,
                            int dtype, void* stream);
/* model_util.py:61-72 + gaze_rnn.py:149-159,381-408: per-frame spatial softmax, soft-label
 * cross-entropy (or l2), and d(loss)/d(logits).  loss_type 0 xentropy, 1 l2, 2 KLD.
 * frame_loss [n_frames] fp32 per-frame partial losses (already divided by B*T);
 * probs / dlogits may be NULL. */
int rgp_softmax_xent_fwd_bwd(const float* logits, const float* gt, float* probs, float* dlogits,
                             float* frame_loss, int n_frames, int hw, float inv_bt, int loss_type,
                             void* stream);
/* deterministic sum of n floats into out[0] (double accumulation) */
int rgp_reduce_sum(const float* x, long long n, float* out, void* stream);

/* ---- optimiser: base.py:262-297, gaze_rnn.py:435-444 ------------------------------------ */
/* sumsq[0] += sum g^2 in double (call rgp_zero first); then clip + TF-Adam in one pass. */
int rgp_sumsq(const float* g, long long n, double* sumsq, void* stream);
/* lr_t = lr*sqrt(1-beta2^t)/(1-beta1^t) is formed on the host in double (TF does the same in
 * its python wrapper); clip <= 0 disables clipping; p_op (may be NULL) receives the operand-dtype copy. */
int rgp_gradnorm_clip_adam(float* p, const float* g, float* m, float* v, long long n,
                           const double* sumsq, float clip, float lr_t, float beta1, float beta2,
                           float eps, void* p_op, int dtype, void* stream);
/* The optimiser schedule evaluated ON THE DEVICE (gaze_rnn.py:439-443 tf.train.exponential_decay(staircase) + the
 * bias-corrected step size tf.train.AdamOptimizer forms in its python wrapper), so that one captured CUDA graph of
 * fwd + BPTT + clip + Adam replays a CORRECT schedule: step[0] (int64, device) = global_step = updates applied so far;
 * lr0[0] (double, device) = initial_learning_rate (the host rewrites it when decay_learning_rate() is called).
 * Writes sched[0] = lr_t = lr*sqrt(1-beta2^t)/(1-beta1^t) with t = step+1, sched[1] = lr = lr0*decay^floor(step/decay_steps),
 * then advances step[0] to t.  rgp_gradnorm_clip_adam_dev is rgp_gradnorm_clip_adam with lr_t read from sched[0]. */
int rgp_adam_schedule(long long* step, const double* lr0, double decay, int decay_steps, double beta1, double beta2,
                      float* sched, void* stream);
int rgp_gradnorm_clip_adam_dev(float* p, const float* g, float* m, float* v, long long n, const double* sumsq,
                               float clip, const float* lr_t_dev, float beta1, float beta2, float eps, void* p_op,
                               int dtype, void* stream);
/* models/base.py:270-273: the reference's other optimisers in TF1's update rules, on the globally clipped gradient
 * (base.py:287-291).  kind 0 = tf.train.MomentumOptimizer(lr, momentum): acc = momentum*acc + g; p -= lr*acc.
 * kind 1 = tf.train.RMSPropOptimizer(lr, decay, momentum, epsilon): ms = decay*ms + (1-decay)*g^2 (ms starts at 1);
 * acc = momentum*acc + lr*g/sqrt(ms+eps); p -= acc.  lr = sched[1] as left by rgp_adam_schedule (which also advances the
 * device-side global_step).  ms may be NULL for kind 0.  p_op: optional bf16 / fp32 operand copy of p (dtype 1 / 0). */
int rgp_gradnorm_clip_momentum_dev(float* p, const float* g, float* acc, float* ms, long long n, const double* sumsq,
                                   float clip, const float* sched, int kind, float momentum, float rms_decay, float eps,
                                   void* p_op, int dtype, void* stream);
int rgp_zero(void* p, long long bytes, void* stream);
/* gaze_rnn.py:504-510 (random h-flip of half the batch) on the fed device buffers: x is [n_clips, rows_per_clip, W]
 * fp32; for each of the nsel clip indices in `clips` (device int32) every W-row is reversed in place. */
int rgp_flip_w(float* x, const int* clips, int nsel, long long rows_per_clip, int W, void* stream);

/* ---- BPTT (base.py:278-281) --------------------------------------------------------------- */
/* Weight-gradient contraction over rows ("wgrad"):
 *   D[tap][i, j] += sum_{m<M} Y[m, ycol0+i] * X[m + tap_off[tap], xcol0+j],  i<Ny, j<Nx
 * Y [>=M, ldy], X [rowsX, ldx] operand dtype row-major; D fp32 [ntaps][Ny][Nx] (the op layout of a
 * conv weight: [tap][Cout][Cin]); accumulates (split-K partials are added atomically). */
typedef struct rgp_wgrad_args {
  const void* Y; int ldy; int ycol0; int Ny;
  const void* X; long long rowsX; int ldx; int xcol0; int Nx;
  int M, ntaps; int tap_off[RGP_MAX_TAPS];
  float* D;
  int dtype, engine;
} rgp_wgrad_args;
int rgp_wgrad(const rgp_wgrad_args* a, void* stream);
/* gate backward of the LAST step (no incoming recurrent gradient): dh = dhs * (gamma*alpha) */
int rgp_gru_bwd_last(const float* dhs, const float* gamma, float alpha, const float* z, const float* c,
                     const float* hprev_f32, void* dg, int ldg, float* dhpart, int B, int G, int Ch,
                     int dtype, void* stream);
/* ConvLSTM gate backward of ONE step (tf.gradients of gaze_lstm.py:112-133 as written).  save [B*S, 5Ch] = the
 * (i, f, g, o, c_{t-1}) rgp_lstm_gates stored for this step; c_t comes from cnext_save (= save of step t+1) or, for
 * the last step, from c_final (p-layout cell buffer).  dhs [B*S, Ch] = head term of dL/dh_t, dhrec [B*S, Ch] = conv^T
 * term from step t+1 (ignored when last != 0); dc [B*S, Ch]: in dL/dc_t (ignored when last), out dL/dc_{t-1}.
 * Outputs (operand dtype, q-layout): dg4 [B*S, 4Ch] = d(pre-activations) i|f|c|o (for W_x and dX), dhc [B*S, 3Ch] =
 * d(hidden conv outputs) (i+c)|f|o (for W_hi, W_hf, W_ho and dh_{t-1}); peephole gradients are accumulated. */
int rgp_lstm_gates_bwd(const float* save, const float* cnext_save, const float* c_final, const float* dhs,
                       const float* dhrec, float* dc, const float* wci, const float* wcf, const float* wco, void* dg4,
                       void* dhc, float* gwci, float* gwcf, float* gwco, int B, int G, int Ch, int last, int dtype,
                       void* stream);
/* per-tap transpose + cast: src fp32 [ntaps][R][Cc] -> dst operand dtype [ntaps][Cc][R] */
int rgp_transpose_taps(const float* src, void* dst, int ntaps, int R, int Cc, int dtype, void* stream);
/* head backward, data side: im2col of dlogits [B,T,49,49] into the head-input row layout:
 * patches[(t*B+b)*100 + (i+3)*10 + j+3, u*19+v] = dlogits[b,t,6i+u-3,6j+v-3] (0 outside), ld = 384 */
int rgp_head_patches(const float* dlogits, void* patches, int B, int T, int dtype, void* stream);
/* head backward, weight side: dW123 [384(361 used), Ch] fp32 -> gradients of w1, w2, w3, out_W (accumulated) */
int rgp_head_compose_bwd(const float* dw123, const float* w1, const float* w2, const float* w3,
                         const float* out_w, const float* w3f, const float* w23, float* dw23, float* dw3f,
                         float* gw1, float* gw2, float* gw3, float* gout_w, int Ch, void* stream);
/* w123 [361, Ch] fp32 -> operand dtype [Ch, 384] (transposed, zero padded) for the dhs GEMM */
int rgp_head_w123t(const float* w123, void* out, int Ch, int dtype, void* stream);
/* frozen-BN backward (gaze_grcn.py:325): dgamma[t,c] += alpha * sum dhs*h ; dbeta[t,c] += sum dhs */
int rgp_bn_bwd(const float* dhs, const float* h_f32, float* dgamma, float* dbeta, float alpha,
               int B, int G, int Ch, void* stream);
/* the same for all T steps in one launch: dhs [T*B*S, Ch] (slab t = step t), h_f32 = slab of h_1 (slab t = h_{t+1}),
 * dgamma / dbeta [T, Ch] */
int rgp_bn_bwd_all(const float* dhs, const float* h_f32, float* dgamma, float* dbeta, float alpha,
                   int B, int T, int G, int Ch, void* stream);
/* out[j] += sum_m X[m, j] (X operand dtype [M, ld]) — bias gradients */
int rgp_colsum(const void* X, int ld, int M, int N, float* out, int dtype, void* stream);
/* gaze_grcn77.py:206-211 backward: dhs[m,c] = dlogit*out_W[c]; gout_W[c] += sum dlogit*h; gout_b += sum dlogit */
int rgp_fc_head_bwd(const float* dlogits, const float* h_f32, const float* out_W, float* dhs,
                    float* gout_W, float* gout_b, int B, int T, int t, int G, int Ch, void* stream);
/* the same for all T steps in one launch: h1_f32 = slab of h_1 (slab t = h_{t+1}), dhs [T*B*S, Ch]; Ch/4 must divide 256, pointers 16-byte aligned */
int rgp_fc_head_bwd_all(const float* dlogits, const float* h1_f32, const float* out_W, float* dhs,
                        float* gout_W, float* gout_b, int B, int T, int G, int Ch, void* stream);

/* ---- dense-GRU variant: models/gaze_rnn.py:283-359 ------------------------------------------
 * Its contractions are plain GEMMs on rgp_tapgemm (ntaps = 1, RGP_EPI_STORE_F32) with widths padded to the engine's
 * granularity (padded weights zero; padded state columns stay zero); these four calls are the element-wise glue.
 * Cell arithmetic = tf.nn.rnn_cell.GRUCell (TF >= 1.2): [r,u] = sig([x,h].Wg + bg); c = tanh([x, r*h].Wc + bc);
 * h' = u*h + (1-u)*c.  Parity unpinned at that library boundary (oracle/model_oracle.py dense_gru_cell). */
/* gaze_rnn.py:296-304,335-338: X[r, pos*Cp + c] = acc[(r*P + pos)*ldacc + c] + bias[c] (flatten order y,x,c), columns
 * P*Cp..ldx-1 zero; X operand dtype [rows, ldx] */
int rgp_dense_pack(const float* acc, int ldacc, const float* bias, void* X, int ldx, long long rows, int P, int Cp,
                   int dtype, void* stream);
/* the same followed by tf.nn.dropout(keep_prob) (gaze_rnn.py:302-303; fed 0.5 while training, :529): kept elements are
 * scaled by 1/keep_prob.  The mask is a counter-based hash of (seed, element index) - reproducible, never stored;
 * rgp_dense_unpack re-derives it.  keep_prob = 1 is rgp_dense_pack. */
int rgp_dense_pack_dropout(const float* acc, int ldacc, const float* bias, void* X, int ldx, long long rows, int P,
                           int Cp, float keep_prob, unsigned long long seed, int dtype, void* stream);
/* xg [M, ldx] fp32 = input halves (columns r | u | c, each Hp wide); hg [M, 2Hp] fp32 = h . Wh_gates; bias_g [2Hp];
 * h_f32 [M, Hp].  Writes rh_op [M, Hp] operand dtype (r*h), u_out [M, Hp] fp32 and (training, else NULL) r_save [M, Hp]. */
int rgp_dense_gru_ru(const float* xg, int ldx, const float* hg, const float* bias_g, const float* h_f32, void* rh_op,
                     float* u_out, float* r_save, int M, int H, int Hp, int dtype, void* stream);
/* hc [M, Hp] fp32 = (r*h) . Wh_cand; bias_c [Hp]; writes h' as fp32 master [M, Hp] and operand copy, and (training, else
 * NULL) the candidate c_save [M, Hp] */
int rgp_dense_gru_c(const float* xg, int ldx, const float* hc, const float* bias_c, const float* u, const float* h_f32,
                    float* hnext_f32, void* hnext_op, float* c_save, int M, int H, int Hp, int dtype, void* stream);
/* ---- BPTT of the dense variant (base.py:278-281 tf.gradients over gaze_rnn.py:283-359) -------------------------
 * dlogits [B, T, N] fp32 -> time-major operand rows dL [(t*B + b), Np] (zero padded) */
int rgp_dense_dlogits(const float* dlogits, void* out, int B, int T, int N, int Np, int dtype, void* stream);
/* one step, first half: dh = dh_head + dh_rec (dh_rec may be NULL: last step) is dL/dh'.  Writes dpre_u = dh (h - c) u (1-u)
 * and dpre_c = dh (1-u)(1-c^2) into dG[:, Hp:2Hp] / dG[:, 2Hp:3Hp] (dG [M, 3Hp] operand dtype) and dhpart = dh u. */
int rgp_dense_gru_bwd_c(const float* dh_head, const float* dh_rec, const float* u, const float* c, const float* hprev,
                        void* dG, float* dhpart, int M, int H, int Hp, int dtype, void* stream);
/* second half, after drh = dpre_c . Wc_h^T: dpre_r = drh h r (1-r) into dG[:, 0:Hp]; dhpart += drh r */
int rgp_dense_gru_bwd_r(const float* drh, const float* r, const float* hprev, void* dG, float* dhpart, int M, int H,
                        int Hp, int dtype, void* stream);
/* backward of rgp_dense_pack[_dropout]: dacc[(r*P + pos), c] = dX[r, pos*Cp + c] (x mask / keep_prob), operand dtype */
int rgp_dense_unpack(const float* dX, int ldx, void* dacc, long long rows, int P, int Cp, float keep_prob,
                     unsigned long long seed, int dtype, void* stream);
/* gaze_rnn.py:346-359: logits[b, t, j] = acc[(t*B + b)*ldacc + j] + bias[j], j < N */
int rgp_dense_logits(const float* acc, int ldacc, const float* bias, float* logits, int B, int T, int N, void* stream);

/* ---- cascade top level: models/gaze_grcn_cascade.py:343-424 ---------------------------
 * Persistent 5x5 ConvGRU with 3 units on the 49x49 grid (:346-351,379; cell arithmetic :95-129): ALL T steps in one
 * launch, one CTA per clip, the state (49*49*3 floats) and r.h live in shared memory for the whole sequence; only the
 * input-half pre-activations are read and the per-step outputs written.  fp32 arithmetic on the CUDA cores for both
 * operand dtypes (P = 2401, Ch = 3: a bandwidth kernel, not a GEMM).
 *   xg    [T*B*So, ldx] fp32: input-half pre-activations (rgp_tapgemm, 25 taps) at q-row f*So + y*Wo + x, f = t*B + b,
 *         columns z(3) | r(3) | c(3);
 *   uzr   [25][3][6] fp32 (tap ky*5+kx, input unit, output z0..2 r0..2) ; uc [25][3][3] fp32;
 *   sal / wsal (both or neither; reading B of the 65-channel declaration, :346-347): sal [B,T,49,49] fp32 frame
 *         saliency, wsal [25][9] fp32 = the 65th input channel's kernels (z|r|c) - its 5x5 conv is added to xg here;
 *   state [B, 7203] fp32 or NULL: in = h_0 when carry != 0 (else h_0 = 0, :353), out = h_T;
 *   og_op [T*B, ldo] operand dtype: row f = h_t flattened (y, x, unit) as tf.reshape(output_g, [B,-1]) does (:384) -
 *         the A operand of the first maxout FC; h_f32 (optional) [T*B, 7203] fp32 the same values. */
int rgp_topgru_seq_fwd(const float* xg, int ldx, int So, int Wo, const float* uzr, const float* uc, const float* sal,
                       const float* wsal, float* state, int carry, void* og_op, int ldo, float* h_f32, int B, int T,
                       int dtype, void* stream);
/* Weight-space composition for RGP_EPI_UP7X: the 11x11 / stride-7 SAME transposed conv (gaze_grcn_cascade.py:317-331)
 * followed by the 5x5 SAME input convs of the top cell (:346-351, z | r | c of 3 units) is linear in the bottom state, so
 *   xg[7A+py, 7B+px, g] = sum_{ky,kx<3} sum_k h[A+ky-1, B+kx-1, k] * Wc[class(A), class(B)][ky, kx][(py*7+px)*16 + g][k],
 *   Wc = sum_{dy,dx in -2..2} sum_c [0 <= 7A+py+dy <= 48][0 <= 7B+px+dx <= 48] wu[7(1-ky)+py+dy+2][7(1-kx)+px+dx+2][c][k]
 *                                                                               * wt[dy+2][dx+2][c][g]
 * (the bracket is the zero border the 5x5 conv sees around the 49x49 map: it only bites in the first / last block row
 * and column, hence the 3 x 3 classes).  up [4][upn][Ch] fp32 = the op layout of the up-sampling filter used with
 * RGP_EPI_UP7 (tap di*2+dj, row (ry*7+rx)*64 + c), wtx [25][16][64] fp32 = the top cell's input kernels (tap, gate, c);
 * wc [9][9][1024][Ch] operand dtype (rows for phases >= 49 / gates >= 9 are written as zeros). */
int rgp_cascade_compose(const float* up, int upn, const float* wtx, void* wc, int Ch, int dtype, void* stream);
/* gaze_grcn_cascade.py:388-420: bias + relu + maxout of one FC layer whose GEMM output y [rows, ldy] fp32 keeps the two
 * halves tf.split(1, 2, .) separates at columns [0, n) and [half, half + n):
 *   v[r, j] = max(relu(y[r, j] + bias[j]), relu(y[r, half + j] + bias[half + j])),  j < n.
 * out_op (operand dtype [rows, ldo], next layer's A operand) and / or out_maps (fp32 [B, T, n]: row r = t*B + b is
 * written to clip b, step t - the predicted gaze maps, :422-424,449) receive v. */
int rgp_maxout(const float* y, int ldy, const float* bias, int half, int n, void* out_op, int ldo, float* out_maps,
               int B, int T, int dtype, void* stream);

/* ---- saliency metrics: evaluation_metrics.py -------------------------------------------- */
/* All metric arithmetic is IEEE float64 with explicit round-to-nearest operations (no FMA contraction):
 * thresholds and integer counts are a bit-exact contract with NumPy's float64 evaluation.
 * Maps are [n, hw] row-major, one frame per row; fix is uint8 (non-zero = fixated). */
/* evaluation_metrics.py:71 (`np.sum(fixation_map > 0.5)` per frame): out[i] = number of non-zero bytes of row i of the
 * uint8 array x[n, row] (fixations per frame, needed on the host before the random index tensors are drawn). */
int rgp_count_nonzero_rows(const uint8_t* x, int n, long long row, int* out, void* stream);

/* evaluation_metrics.py:15-17,246: fp32 model output [n,hw] -> float64, optionally (x-min)/(max-min) per frame */
int rgp_metrics_prep(const float* src, double* dst, int n, int hw, int normalize, void* stream);
/* stand-in for skimage.transform.resize(order=3) (evaluation_metrics.py:76,137,248,255; see
 * oracle/metrics_oracle.py resize_bicubic): [n,h,w] -> [n,oh,ow]; mode_nearest 0 = 'constant', 1 = 'nearest' */
int rgp_resize_bicubic(const double* src, double* dst, int n, int h, int w, int oh, int ow,
                       int mode_nearest, void* stream);
/* evaluation_metrics.py:207-236 (+NSS): out [n,8] = sim, cc, nss, min(pred), max(pred), n_fix, sum(pred), sum(gt).
 * fix may be NULL (nss = NaN). */
int rgp_sim_cc_nss(const double* pred, const double* gt, const uint8_t* fix, double* out,
                   int n, int hw, void* stream);
/* evaluation_metrics.py:42-98: AUC-Judd per frame.  sal [n,hw] f64 (resized prediction), jitter [n,hw] f64 = the
 * host's random.rand(H,W) draw (NULL: jitter=False).  Outputs: thr [n,maxfix] f64 = normalised fixation values
 * sorted descending (NaN padded), counts [n,maxfix] int64 = #{S >= thr_k} (-1 padded; bit-exact contract),
 * auc [n] f64 (NaN: no fixation; -1: more than maxfix fixations).  maxfix <= 1024. */
int rgp_auc_judd(const double* sal, const double* jitter, const uint8_t* fix, double* thr, long long* counts,
                 double* auc, int n, int hw, int maxfix, void* stream);
/* evaluation_metrics.py:101-204: AUC-Borji / shuffled AUC with the host-drawn index tensor idx [n,nrows,nrep]
 * int32 (pixel index of each negative; -1 = absent row).  sal [n,hw] f64 un-normalised, with per-frame
 * (min,max) at minmax[f*mm_stride + {0,1}] (e.g. columns 3,4 of rgp_sim_cc_nss's output).  auc [n] f64 = mean over
 * repetitions; tpc/fpc (optional, both or neither) int32 [n,nrep,nthr_max]: counts at threshold i*step. */
int rgp_auc_sampled(const double* sal, const uint8_t* fix, const int* idx, const double* minmax, int mm_stride,
                    double* auc, int* tpc, int* fpc, int n, int hw, int nrows, int nrep, double step,
                    int nthr_max, int maxfix, void* stream);

/* ---- fused resize + metric (evaluation_metrics.py:246-257 followed by :42-236) --------------
 * The same three metric kernels with the bicubic enlargement of the model-resolution map(s) [n,h,w] to the fixation
 * map's size [oh,ow] evaluated on the fly (never written to HBM).  Every resized value is bit-identical to
 * rgp_resize_bicubic's, so thresholds / integer counts / AUCs equal the two-step path exactly; the sums of
 * sim / cc / NSS are taken strip-major (differences ~1e-16).  fix / jitter / idx / outputs as in the plain calls. */
/* 1 if the strip tables of an h x w -> oh x ow enlargement fit in shared memory (else use the two-step path) */
int rgp_metrics_fused_fits(int h, int w, int oh, int ow);
int rgp_sim_cc_nss_resized(const double* pred_src, const double* gt_src, const uint8_t* fix, double* out,
                           int n, int h, int w, int oh, int ow, int mode_nearest, void* stream);
int rgp_auc_judd_resized(const double* sal_src, const double* jitter, const uint8_t* fix, double* thr,
                         long long* counts, double* auc, int n, int h, int w, int oh, int ow, int mode_nearest,
                         int maxfix, void* stream);
/* minmax: per-frame (min,max) of the RESIZED map (columns 3,4 of rgp_sim_cc_nss_resized's output) */
int rgp_auc_sampled_resized(const double* sal_src, const uint8_t* fix, const int* idx, const double* minmax,
                            int mm_stride, double* auc, int* tpc, int* fpc, int n, int h, int w, int oh, int ow,
                            int mode_nearest, int nrows, int nrep, double step, int nthr_max, int maxfix,
                            void* stream);

/* ---- opt-in device random numbers for the evaluation (csrc/rng.cu): statistically equivalent to, NOT stream-identical with,
 * the NumPy global-RNG draws of evaluation_metrics.py:78-79 (jitter) and :148 (Borji negatives) ------------------------ */
/* out[0..n) = uniform doubles in [0,1), a counter-based hash of (seed, index); out 16-byte aligned */
int rgp_fill_uniform_f64(double* out, long long n, unsigned long long seed, void* stream);
/* out [nframes, nrows, nrep] int32: rows r < nfix[f] uniform pixels in [0, npix), rows r >= nfix[f] = -1 (absent) */
int rgp_fill_randint_rows(int* out, const int* nfix, int nframes, int nrows, int nrep, int npix, unsigned long long seed,
                          void* stream);

const char* rgp_last_error(void);
int rgp_version(void);
/* 1 if the tcgen05 engine can run on the current device (sm_100), else 0 */
int rgp_has_tcgen05(void);

#ifdef __cplusplus
}
#endif
#endif
