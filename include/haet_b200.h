/*
 * haet_b200.h -- C-ABI of the B200-native HAET layer hot path (libhaet_b200.so).
 *
 * Plain pointers and sizes only: no torch types.  Every pointer named *_d is a
 * DEVICE pointer on the current CUDA device; `stream` is a cudaStream_t passed
 * as void*.  All entry points only enqueue work on `stream` (no host sync, CUDA
 * graph capturable) and return 0 on success or a negative HAET_E* code;
 * haet_last_error() gives the message.  The caller owns every buffer; inputs
 * are never written.  There is no CPU fallback: without an sm_100 device the
 * calls fail with HAET_ENODEV.
 *
 * Which reference interface each entry point replaces (paths relative to
 * /root/reference, the pedrocurvo/HAET checkout):
 *
 *   haet_slice_reduce      models/PhysicsAttention/IrregularMesh.py:157-188
 *                          (in_project_x, Ada-Temp, Rep-Slice softmax over the
 *                          N points, W^T x), conv front-ends of
 *                          StructuredMesh2D.py:173-189 / StructuredMesh3D.py:177-193
 *   haet_slice_merge       IrregularMesh.py:177-209 (mass normalisation, eidetic
 *                          states, slice positions) + the cross-rank merge of the
 *                          point-sharded multi-GPU layout (new)
 *   haet_balltree_build    models/components/balltree/balltree.pyx:598-662
 *                          build_balltree_with_rotations (and balltree.h)
 *   haet_erwin_forward     models/components/erwinflash/erwin_flash.py:138-195
 *                          (+ components/{layers,attention,mlp,pooling}.py)
 *   haet_deslice_out       IrregularMesh.py:221-235 (deslice W.S' + to_out)
 *   haet_slice_reduce_ln / haet_deslice_out_res
 *                          the same with the caller's ln_1 and residual add fused in
 *                          (models/HAETransolver_Irregular_Mesh.py:125)
 *   haet_model_prologue / haet_time_embed / haet_head_ln_linear
 *                          Model.forward around the blocks: preprocess MLP + placeholder + time embedding
 *                          (HAETransolver_Irregular_Mesh.py:351-370, 297-336) and the last block's mlp2(ln_3(fx)) (:131-132)
 *   haet_slice_reduce_halo   haet_slice_reduce for one slab of a structured mesh sharded over GPUs (new)
 *   haet_layer_forward_host  the whole Physics_Attention_*.forward
 *                          (IrregularMesh.py:141-235) on HOST buffers
 */
#ifndef HAET_B200_H
#define HAET_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HAET_OK 0
#define HAET_EINVAL (-1)  /* bad argument (the reference raises ValueError / AssertionError) */
#define HAET_ECUDA (-2)   /* a CUDA call failed */
#define HAET_ENODEV (-3)  /* no sm_100 device */
#define HAET_ENOSUP (-4)  /* shape outside what the kernels support */

#define HAET_MAX_LEVELS 8

/* Shape of one PhysicsAttention layer (IrregularMesh.py:41-63 constructor arguments). */
typedef struct haet_layer_shape {
    int64_t B;          /* batch */
    int64_t N;          /* points per batch item (H*W[*D] for structured meshes) */
    int32_t dim;        /* input/output channels (n_hidden) */
    int32_t heads;      /* H */
    int32_t dim_head;   /* C; inner_dim = heads*dim_head */
    int32_t slice_num;  /* G */
    int32_t conv_nd;    /* 0: Linear in_project_x; 2: Conv2d; 3: Conv3d (kernel k, pad k/2) */
    int32_t grid[3];    /* structured grid extents (H, W, D); unused when conv_nd == 0 */
    int32_t kernel;     /* conv kernel size (3) */
    int32_t dimensionality; /* 3 irregular/3D, 2 for 2D: #channels used as slice positions */
    float base_temp;    /* 0.5 */
    float epsilon;      /* 1e-6 */
} haet_layer_shape;

/* Hyper-parameters of the Erwin transformer over the slice tokens
 * (erwin_flash.py:42-58 constructor arguments; embed=False, mp_steps=0). */
typedef struct haet_erwin_cfg {
    int32_t n_levels;                    /* len(ball_sizes) = encoder levels + bottleneck */
    int32_t c_hidden[HAET_MAX_LEVELS];
    int32_t ball_sizes[HAET_MAX_LEVELS];
    int32_t enc_num_heads[HAET_MAX_LEVELS];
    int32_t enc_depths[HAET_MAX_LEVELS];
    int32_t dec_num_heads[HAET_MAX_LEVELS]; /* n_levels-1 entries */
    int32_t dec_depths[HAET_MAX_LEVELS];
    int32_t strides[HAET_MAX_LEVELS];
    float rotate;                        /* rotation angle in degrees; > 0 enables the rotated permutation */
    int32_t decode;
    int32_t mlp_ratio;
    int32_t dimensionality;
    int32_t attn_bias;                   /* 0: erwinflash BallMSA (no bias, the HAET default); 1: add the distance bias
                                          * sigma_att[h] * ||p_i - p_j|| of the non-flash copy (benchmarks/02-.../erwin/components/
                                          * attention.py:28-32,49-51); every block's packed weights then end with sigma_att[heads] */
    int32_t bn_train;                    /* 0: eval-mode BatchNorm1d of BallPooling / BallUnpooling (pooling.py:23,56) folded into a per-channel
                                          * (scale, shift) by the caller.  1: module.training semantics (forward only): the
                                          * (scale, shift) slots of the packed weights hold (gamma, beta), every BatchNorm normalises with
                                          * the batch mean / biased variance over ALL rows of its level (duplicated leaf slots included,
                                          * as in the reference) */
    float bn_eps;                        /* BatchNorm1d.eps (1e-5), used when bn_train */
    float *bn_stats_d;                   /* bn_train: NULL, or device floats receiving (mean[C], biased var[C]) of every BatchNorm in
                                          * execution order: encoder pools 0 .. L-1 (strides != 1), then decoder unpools L-1 .. 0 */
} haet_erwin_cfg;

int haet_version(void);
const char *haet_last_error(void);
/* 0 when the current device is sm_100 and the kernels can run, else HAET_ENODEV. */
int haet_device_check(void);

/* ---- sizes of the caller-allocated buffers ---------------------------------------- */
/* bytes of the opaque per-point spill written by haet_slice_reduce and read by haet_deslice_out */
int64_t haet_spill_bytes(const haet_layer_shape *s);
/* bytes of scratch haet_slice_reduce / haet_deslice_out need */
int64_t haet_slice_workspace_bytes(const haet_layer_shape *s);
int64_t haet_deslice_workspace_bytes(const haet_layer_shape *s);
/* floats in one packed rank-partial slice state: B*heads*G*(C+2)  [(m2, l, S[C]) per slice] */
int64_t haet_partial_floats(const haet_layer_shape *s);
/* number of floats of the packed Erwin weights / of its scratch, for T input tokens */
int64_t haet_erwin_weight_floats(const haet_erwin_cfg *c);
int64_t haet_erwin_workspace_bytes(const haet_erwin_cfg *c, int64_t n_slots);
/* slots of the main ball tree for n_trees trees of pts points each (2^ceil(log2 pts) each) */
int64_t haet_balltree_slots(int64_t n_trees, int64_t pts);

/* ---- slicing (pass 1 over the points) ---------------------------------------------
 * x_d      [B, N, dim] fp32
 * wx_d     Linear: [inner, dim];  conv: [inner, taps, dim] (taps = k^nd, tap-major repack of
 *          the torch [inner, dim, k, k(, k)] weight -- done once by the host module)
 * ws_d     in_project_slice.weight [G, C];  bs_d [G];  wt_d ada_temp_linear.weight [C]; bt scalar
 * spill_d  haet_spill_bytes() bytes (projected points kept for the deslice pass)
 * partial_d haet_partial_floats() floats: this rank's un-normalised slice state in the
 *          base-2 log domain: m2 = a reference exponent (<= max_n z*log2e), l = sum_n 2^(z2-m2), S = sum_n 2^(z2-m2) x_proj
 * dim_head: 32 and 16 (zero-padded image rows) and 64 (two 32-channel pseudo-heads per head, slice_num <= 256) run on the
 * tcgen05 passes; 4 / 8, and 64 with more slices, on the fp32 SIMT kernels.
 * Numeric range: the tensor-core path keeps x_proj as fp16 hi + lo (relative 2^-22 for |x_proj| in [2^-3, 65504]); values
 * beyond +-65504 SATURATE (they never become inf / NaN) and values below 2^-14 keep an absolute precision of 2^-25.
 */
int haet_slice_reduce(const haet_layer_shape *s, const float *x_d, const float *wx_d, const float *bx_d,
                      const float *ws_d, const float *bs_d, const float *wt_d, float bt,
                      void *spill_d, float *partial_d, void *workspace_d, void *stream);

/* Block fusion, SURVEY 8f rank 1 (models/HAETransolver_Irregular_Mesh.py:125  fx = self.Attn(self.ln_1(fx)) + fx):
 * haet_slice_reduce_ln == haet_slice_reduce(LayerNorm(x; ln_g_d, ln_b_d, ln_eps)) with x_d the RAW residual stream.  On the
 * tensor-core Linear front-end (heads*dim_head == 256) the in-projection kernel computes the row statistics while it
 * converts x and applies them in its epilogue (weights folded on the device: no normalised copy of x ever exists); other
 * shapes normalise into the workspace first.  workspace_d must hold haet_slice_ln_workspace_bytes() bytes. */
int64_t haet_slice_ln_workspace_bytes(const haet_layer_shape *s);
int haet_slice_reduce_ln(const haet_layer_shape *s, const float *x_d, const float *ln_g_d, const float *ln_b_d, float ln_eps,
                         const float *wx_d, const float *bx_d, const float *ws_d, const float *bs_d, const float *wt_d,
                         float bt, void *spill_d, float *partial_d, void *workspace_d, void *stream);

/* ---- slab sharding of the structured meshes (new; SURVEY 8e "2D/3D" + 8f-4 halo exchange) --------------------------
 * A rank owns grid[0] consecutive planes of the (H, W[, D]) grid -- a contiguous run of rows of x -- and the conv front-end
 * (StructuredMesh2D.py:173-189, StructuredMesh3D.py:177-193) needs kernel / 2 planes of its neighbours on either side.
 * s describes the LOCAL slab (grid[0] = local planes, N = local points); x_ext_d [B, (halo_lo + grid[0] + halo_hi) * plane, dim]
 * holds the neighbour's planes, then the slab, then the other neighbour's planes (the caller exchanges them: haet_b200/
 * distributed.py: exchange_halo).  halo_lo / halo_hi are 0 at an outer boundary of the grid (the conv's zero padding is then the
 * true padding) and kernel / 2 otherwise.  ln_g_d / ln_b_d (both NULL or both set): the caller's ln_1 applied to every row first.
 * Outputs are those of haet_slice_reduce for the slab's own points; everything downstream (haet_peer_exchange / all-gather,
 * haet_slice_merge with n_ranks = world, ball tree, Erwin, haet_deslice_out on the slab's spill) is the point-sharded flow of the
 * irregular model.  Tensor-core passes only (dim_head 16 / 32 / 64; else HAET_ENOSUP). */
int64_t haet_slice_halo_workspace_bytes(const haet_layer_shape *s, int32_t halo_lo, int32_t halo_hi, int32_t with_ln);
int haet_slice_reduce_halo(const haet_layer_shape *s, int32_t halo_lo, int32_t halo_hi, const float *x_ext_d, const float *ln_g_d,
                           const float *ln_b_d, float ln_eps, const float *wx_d, const float *bx_d, const float *ws_d,
                           const float *bs_d, const float *wt_d, float bt, void *spill_d, float *partial_d, void *workspace_d,
                           void *stream);

/* Merge R packed rank-partials (rank-major, fixed order => bit-reproducible on every rank) into
 *   eidetic_d [B*heads*G, C]  = (S/l)/(1+1e-5)          (IrregularMesh.py:181-188; sum_n W == 1)
 *   lse2_d    [B*heads*G]     = m2 + log2(l)            (global softmax normaliser, base 2)
 *   pos_d     [B*heads*G, dimensionality]               (IrregularMesh.py:204-209)
 */
int haet_slice_merge(const haet_layer_shape *s, const float *partials_d, int32_t n_ranks,
                     float *eidetic_d, float *lse2_d, float *pos_d, void *stream);

/* ---- cross-GPU exchange of the packed slice state over peer memory (new) ----
 * Replaces the all-gather between the two merge levels by a kernel working on NVLink peer memory (csrc/peer_exchange.cu).
 * peer_bufs: HOST array of `world` device pointers, entry r = rank r's symmetric buffer of haet_peer_buffer_bytes() bytes as
 * mapped in this process (zero-initialised once; e.g. torch.distributed._symmetric_memory); partial_d: this rank's packed
 * state from haet_slice_reduce; gathered_d [world * haet_partial_floats()] is what haet_slice_merge(n_ranks = world) reads;
 * state_d: 4 zero-initialised uint32 in local memory (device-side epoch, tickets, error word).  No host-side counter:
 * graph capturable.  Every rank must make the same sequence of calls. */
int64_t haet_peer_buffer_bytes(const haet_layer_shape *s);
int haet_peer_exchange(const haet_layer_shape *s, void *const *peer_bufs, int32_t rank, int32_t world,
                       const float *partial_d, float *gathered_d, void *state_d, void *stream);

/* ---- ball tree (balltree.pyx:598-662) ----------------------------------------------
 * pos_d [n_trees*pts, d] fp32; every tree has `pts` >= 2 consecutive rows (batch_idx = row / pts).
 * tree_idx_d int64[slots], tree_mask_d uint8[slots]; rot_idx_d[i] int64[slots / prod(strides[:i])]
 * for i < n_levels (ignored when angle_deg <= 0); rot_idx_d itself is a HOST array of n_levels
 * device pointers.  Bit-exact with the reference extension,
 * including the element order std::nth_element leaves behind. */
int haet_balltree_build(const float *pos_d, int64_t n_trees, int64_t pts, int32_t d,
                        const int32_t *strides, const int32_t *ball_sizes, int32_t n_levels,
                        double angle_deg, int64_t *tree_idx_d, uint8_t *tree_mask_d,
                        int64_t *const *rot_idx_d, void *stream);

/* ---- Erwin over the slice tokens (erwin_flash.py:138-195, eval-mode BatchNorm) -------
 * x_d [T, c_hidden[0]], pos_d [T, d]; tree from haet_balltree_build (n_slots entries);
 * weights_d: packed fp32 parameters in the order documented in haet_b200/erwin.py (pack order) / csrc/erwin.cu (Walker);
 * out_d [T, c_hidden[0]] in the original token order (erwin_flash.py:193). */
int haet_erwin_forward(const haet_erwin_cfg *c, const float *x_d, const float *pos_d, int64_t T,
                       const int64_t *tree_idx_d, const uint8_t *tree_mask_d, int64_t n_slots,
                       int64_t *const *rot_idx_d, const float *weights_d, void *workspace_d,
                       float *out_d, void *stream);

/* The same with the number of trees given (n_slots / n_trees slots each).  Default: one launch per block (programmatic dependent
 * launch).  With HAET_ERWIN_FUSED=1 the pass runs as ONE cooperative kernel whose phases (gather, blocks, pool / unpool, scatter)
 * are separated by per-TREE barriers -- nothing in Erwin crosses a tree; bitwise the same result, measured slower.
 * haet_erwin_forward is this call with n_trees = 1 (one barrier domain). */
int haet_erwin_forward_trees(const haet_erwin_cfg *c, const float *x_d, const float *pos_d, int64_t T, int64_t n_trees,
                             const int64_t *tree_idx_d, const uint8_t *tree_mask_d, int64_t n_slots,
                             int64_t *const *rot_idx_d, const float *weights_d, void *workspace_d,
                             float *out_d, void *stream);

/* haet_erwin_forward_trees WITHOUT a caller workspace.  No counterpart in the reference: added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  The scratch is a file-scope device array of HAET_ERWIN_STATIC_SCRATCH_BYTES inside the library (csrc/erwin.cu):
 * every GPU that loads the library holds it until exit, and the call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine).  Inputs and outputs are still the caller's.  Fails with HAET_ENOSUP -- and launches nothing --
 * when haet_erwin_workspace_bytes(c, n_slots) exceeds the array (the default HAET configuration, c_hidden {C, 2C} with C = 32,
 * takes 274 bytes per slot: ~30,000 slots = B*heads*G slice tokens; the layer's own sizes are 512..4096).
 * EVERY call on a device shares the one array, which keeps the previous call's contents (the call clears the part that must
 * start at zero): two calls that overlap on the GPU -- different streams, or graphs replayed concurrently -- RACE; serialise
 * them, or use haet_erwin_forward_trees with one workspace per stream.  The array's address is looked up on the first call on
 * each device (make one un-captured call per device before capturing a graph). */
#define HAET_ERWIN_STATIC_SCRATCH_BYTES (8 * 1024 * 1024)
int haet_erwin_forward_static(const haet_erwin_cfg *c, const float *x_d, const float *pos_d, int64_t T, int64_t n_trees,
                              const int64_t *tree_idx_d, const uint8_t *tree_mask_d, int64_t n_slots,
                              int64_t *const *rot_idx_d, const float *weights_d, float *out_d, void *stream);

/* ---- deslice + to_out (pass 2 over the points) -------------------------------------
 * states_d [B*heads*G, C] = Erwin output; lse2_d from haet_slice_merge; out_d [B, N, dim]. */
int haet_deslice_out(const haet_layer_shape *s, const void *spill_d, const float *lse2_d,
                     const float *states_d, const float *ws_d, const float *bs_d, const float *wt_d,
                     float bt, const float *wo_d, const float *bo_d, float *out_d,
                     void *workspace_d, void *stream);

/* Same with the block's residual connection fused into the to_out epilogue: out = deslice_out(...) + resid_d
 * (resid_d [B, N, dim], may be NULL, must not alias out_d). */
int haet_deslice_out_res(const haet_layer_shape *s, const void *spill_d, const float *lse2_d, const float *states_d,
                         const float *ws_d, const float *bs_d, const float *wt_d, float bt, const float *wo_d,
                         const float *bo_d, const float *resid_d, float *out_d, void *workspace_d, void *stream);

/* ---- the MLP half of the block (new) ------------------------------------------
 * out = W2 . gelu(W1 . LayerNorm(x) + b1) + b2 + x
 * replaces  fx = self.mlp(self.ln_2(fx)) + fx  (models/HAETransolver_Irregular_Mesh.py:128) with
 * MLP(n_layers=0, act=gelu) = Linear(dim, hidden) -> nn.GELU() -> Linear(hidden, dim)  (models/components/mlp.py:67-96).
 * x_d / out_d [M, dim] fp32 (must not alias), w1_d [hidden, dim], w2_d [dim, hidden].  Two tensor-core GEMM passes: ln_2 is
 * fused into the first (row statistics in its producers), GELU into its epilogue, which writes the hidden activations as
 * the second GEMM's bf16 hi/lo operand image; bias and residual ride in the second epilogue.  Supported: dim == 256 and
 * hidden a multiple of 256 (else HAET_ENOSUP; haet_block_mlp_workspace_bytes returns HAET_ENOSUP as well). */
int64_t haet_block_mlp_workspace_bytes(int64_t M, int32_t dim, int32_t hidden);
int haet_block_mlp(const float *x_d, int64_t M, int32_t dim, int32_t hidden, const float *ln_g_d, const float *ln_b_d,
                   float ln_eps, const float *w1_d, const float *b1_d, const float *w2_d, const float *b2_d, float *out_d,
                   void *workspace_d, void *stream);

/* ---- model prologue / epilogue around the blocks (new) ----------------------------
 * haet_model_prologue replaces, in Model.forward (models/HAETransolver_Irregular_Mesh.py:351-370; the structured models'
 * HAETransolver_Structured_Mesh_2D.py:384-410 is the same computation),
 *     fx = preprocess(cat(x, fx)) + placeholder [+ time_fc(timestep_embedding(T))]
 * with preprocess = MLP(n_layers = 0, act = gelu) = Linear(K, hidden) -> GELU -> Linear(hidden, dim)  (components/mlp.py:67-96):
 *     out[m, :] = W2 . gelu(W1 . [feat_a[m] | feat_f[m]] + b1) + b2 + add
 * feat_a_d [B*N, Ka] (the coordinates, or precomputed position features) and feat_f_d [B*N, Kf] (the function values) are
 * concatenated on the fly; either may be NULL with its width 0.  ref > 0 (unified_pos of the irregular model, get_grid,
 * :297-336): feat_a_d holds 2-D coordinates (Ka == 2) and the features are their ref*ref distances to the reference grid.
 * add_d: NULL, or [dim] (placeholder; add_per_batch = 0), or [B, dim] (placeholder + time embedding from haet_time_embed;
 * add_per_batch = 1).  Supported: dim == 256, hidden a multiple of 256 (<= 2048), Ka(ref^2) + Kf <= 128; otherwise HAET_ENOSUP
 * (the caller then runs its own modules).  The hidden activations only exist as the second GEMM's operand image. */
int64_t haet_model_prologue_workspace_bytes(int64_t M, int32_t Ka, int32_t Kf, int32_t ref, int32_t hidden, int32_t dim);
int haet_model_prologue(const float *feat_a_d, int32_t Ka, const float *feat_f_d, int32_t Kf, int32_t ref, int64_t B, int64_t N,
                        int32_t hidden, int32_t dim, const float *w1_d, const float *b1_d, const float *w2_d, const float *b2_d,
                        const float *add_d, int32_t add_per_batch, float *out_d, void *workspace_d, void *stream);
/* add_d[b, :] = base_d[:] (may be NULL) + W2 . silu(W1 . timestep_embedding(t_d[b], dim) + b1) + b2
 * (models/components/embedding.py:72-99; time_fc of HAETransolver_Irregular_Mesh.py:220-223, :367-370) */
int haet_time_embed(const float *t_d, int64_t B, int32_t dim, const float *w1_d, const float *b1_d, const float *w2_d,
                    const float *b2_d, const float *base_d, float *add_d, void *stream);
/* y[m, :] = W . LayerNorm(x[m, :]) + b  -- the last block's mlp2(ln_3(fx)) (HAETransolver_Irregular_Mesh.py:131-132) in one
 * pass over x.  dim a multiple of 128 (<= 1024), out_dim 1..8, else HAET_ENOSUP. */
int haet_head_ln_linear(const float *x_d, int64_t M, int32_t dim, const float *ln_g_d, const float *ln_b_d, float ln_eps,
                        const float *w_d, const float *b_d, int32_t out_dim, float *y_d, void *stream);

/* ---- whole layer on HOST buffers (copies inside) ------------------------------------
 * The call a non-torch host (C, C++, ctypes) makes: x_h / out_h are host pointers (pinned for
 * full speed); weights are device pointers prepared once.  Allocates nothing: scratch_d must
 * hold haet_layer_scratch_bytes() bytes. */
typedef struct haet_layer_weights {
    const float *wx_d, *bx_d, *ws_d, *bs_d, *wt_d, *wo_d, *bo_d, *erwin_d;
    float bt;
} haet_layer_weights;
int64_t haet_layer_scratch_bytes(const haet_layer_shape *s, const haet_erwin_cfg *c);
int haet_layer_forward_host(const haet_layer_shape *s, const haet_erwin_cfg *c, const haet_layer_weights *w,
                            const float *x_h, float *out_h, void *scratch_d, void *stream);
/* Same with device x/out (what the torch module calls when it does not need the stages). */
int haet_layer_forward(const haet_layer_shape *s, const haet_erwin_cfg *c, const haet_layer_weights *w,
                       const float *x_d, float *out_d, void *scratch_d, void *stream);

/* ---- switches and test hooks -------------------------------------------------------
 * The three setters below are THREAD-LOCAL: they change what later calls made by the SAME host thread launch and
 * nothing else.  The library keeps no other mutable state (the ablation switches of INTEGRATION.md are read from the
 * environment once, at first use), so the entry points are re-entrant across threads and streams.
 * haet_set_path(0): tcgen05 tensor-core kernels wherever the shape allows (default);
 * haet_set_path(1): the exact-fp32 SIMT CUDA kernels everywhere (device-side cross-check).
 * haet_debug_gemm: Y[M,Nn] = X[M,K].W[Nn,K]^T + bias through the path selected above;
 * img_d: scratch of 4*Nn*K bytes. */
int haet_set_path(int path);
/* bench hook: which kernels haet_slice_reduce / haet_deslice_out launch (default 15 = all), so that each can be timed
 * alone through the real entry points on buffers a full call has filled:
 * bit 0 in-projection (+ image pack), bit 1 pass-1 slice kernel + chunk merge, bit 2 pass-2 deslice kernel, bit 3 to_out */
int haet_set_stage_mask(int mask);
/* test hook: lazy-max threshold of pass 1 in log2 units, (0, 15]; 0 restores the default (15).  A small value forces the
 * accumulator-rescale path on ordinary inputs. */
int haet_set_raise_threshold(float log2_units);
int haet_debug_gemm(const float *x_d, const float *w_d, const float *bias_d, float *y_d, int64_t M, int32_t K,
                    int32_t Nn, void *img_d, void *stream);

/* number of kernels launched by this library since load (bench.py's gpu_launches) */
int64_t haet_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* HAET_B200_H */
