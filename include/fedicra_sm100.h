/*
 * fedicra_sm100.h -- C ABI of libfedicra_sm100.so, the B200 (sm_100a) implementation of FedICRA's weak-supervision
 * loss hot path.  This is the drop-in boundary: plain pointers and sizes, no torch types.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer on the current CUDA device unless its name starts with `h_`;
 *   - tensors are dense, row-major ("contiguous"), float = IEEE fp32, int = int32;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), never synchronises the host and
 *     never allocates: scratch memory comes from the caller (`workspace`, size from the matching *_workspace_bytes);
 *   - return value 0 = launched; negative = rejected (bad argument / launch failure), text in fedicra_last_error();
 *   - there is no CPU fallback anywhere behind these symbols.
 *
 * Reference interfaces replaced (paths relative to /root/reference/code, LTF = utils/TreeEnergyLoss/kernels/
 * lib_tree_filter):
 *   pybind module tree_filter_cuda                       LTF/src/tree_filter.cpp:7-13
 *   ModelLossSemsegGatedCRF.forward                      utils/gate_crf_loss.py:20-124
 *   TreeFilter2D.build_edge_weight                       LTF/modules/tree_filter.py:91-108
 *   MinimumSpanningTree._build_feature_weight            LTF/modules/tree_filter.py:27-34
 *   MyModel.set_weights element-wise update              flower_common.py:553-555, 590-602
 */
#ifndef FEDICRA_SM100_H
#define FEDICRA_SM100_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Text of the last error raised on the calling thread ("" if none). */
const char *fedicra_last_error(void);
/* ABI version of this header (bumped on any signature change or added entry point; 4 = round 2: chain order / chain
 * filter, armed filter buffers, bilinear resize, MST profile). */
int fedicra_abi_version(void);
/* Telemetry: kernels launched by this library in this process so far (no counterpart in the reference). */
long long fedicra_launch_count(void);

/* ------------------------------------------------------------------------------------------------ Gated CRF ---- */
/* One kernel descriptor of the Gaussian mixture: K += weight * exp(-0.5*(|dxy|^2/sigma_xy^2 + |dI|^2/sigma_rgb^2)).
 * A sigma <= 0 means "modality not used by this descriptor".   (gate_crf_loss.py:134-161, kernels_desc entries) */
typedef struct {
    float weight;
    float sigma_xy;
    float sigma_rgb;
} fedicra_gcrf_desc;

/* Scratch bytes needed by fedicra_gcrf_fwd_bwd for an [N,*,H,W] problem.  The scratch must be zero-filled once
 * before its first use; the kernel leaves it reusable. */
size_t fedicra_gcrf_workspace_bytes(int N, int H, int W);

/* Fused forward + backward of ModelLossSemsegGatedCRF.forward (gate_crf_loss.py:20-124, live branch: no masks,
 * Potts compatibility), without materialising any unfold tensor (replaces :89-99 and :163-172).
 *   y        [N,C,H,W]   softmax probabilities
 *   image    [N,Cin,H,W] the rgb modality already at prediction resolution (pooling, :126-132, stays on the host side)
 *   loss     [1]         (sum K - sum_c y*(K (*) y)) / (N*H*W)
 *   grad_y   [N,C,H,W]   d loss / d y  = -2 (K (*) y) / (N*H*W)            (may be NULL: forward only)
 * Out-of-image taps see zero features, exactly like F.unfold's zero padding (:183-188). */
int fedicra_gcrf_fwd_bwd(const float *y, const float *image, int N, int C, int Cin, int H, int W, int radius,
                         const fedicra_gcrf_desc *h_descs, int n_desc, float *loss, float *grad_y, void *workspace,
                         size_t workspace_bytes, void *stream);

/* Allocation-free variant of fedicra_gcrf_fwd_bwd.  It has no counterpart in the reference.
 * For callers that cannot create device memory themselves (their host framework owns all of it, or they run where a
 * cudaMalloc -- which synchronises the device -- is not acceptable): the scratch is a fixed array inside the library,
 * fedicra_gcrf_noalloc_max_bytes() = 131136 bytes per GPU, enough for 32 768 thread blocks: max(N*ceil(H/16)*ceil(W/32), ceil(N*H*W/128)) <= 32768, i.e. N*H*W <= 4 194 304
 * pixels (16 x 512 x 512) for ordinary image shapes.
 * Same arguments and results as fedicra_gcrf_fwd_bwd minus the workspace; problems beyond the limit are refused with an
 * error.  The array is shared: calls that can overlap on the GPU race, serialise them.  The first call on a device
 * performs a one-off address lookup and should be made before any stream capture on that device. */
size_t fedicra_gcrf_noalloc_max_bytes(void);
int fedicra_gcrf_fwd_bwd_noalloc(const float *y, const float *image, int N, int C, int Cin, int H, int W, int radius,
                                 const fedicra_gcrf_desc *h_descs, int n_desc, float *loss, float *grad_y,
                                 void *stream);

/* ------------------------------------------------------------------------------------------------ MST ---------- */
/* Scratch bytes for the two MST entry points (B images, V vertices, E candidate edges each). */
size_t fedicra_mst_workspace_bytes(int B, int V, int E);

/* mst_forward (LTF/src/mst/mst.cu:86-117 + boruvka.cpp:20-112): minimum spanning tree of B graphs given as explicit
 * edge lists.  Ties are broken by edge index, which is the edge set the reference's sequential Boruvka produces.
 *   edge_index  [B,E,2] int32, edge_weight [B,E] fp32 (no NaN)  ->  edge_out [B,V-1,2] int32 in increasing edge index.
 * Unlike the reference it never leaves the device (no D2H/H2D, no host threads). */
int fedicra_mst_forward(const int *edge_index, const float *edge_weight, int B, int E, int V, int *edge_out,
                        void *workspace, size_t workspace_bytes, void *stream);

/* MinimumSpanningTree.forward on the implicit 4-connected grid (LTF/modules/tree_filter.py:15-34, 53-62 fused with
 * mst_forward): edge e<(H-1)*W joins (h,w)-(h+1,w), the rest join (h,w)-(h,w+1); weight = sum_c (delta)^2 + 1 summed in
 * channel order without fused multiply-add, so the fp32 bits equal torch's for D <= 4 (what FedICRA uses: 2..3; torch's
 * CUDA reduction interleaves four accumulators beyond that -- the Python modules take torch's own weights for D > 4).
 *   embed [B,D,H,W]  ->  edge_out [B,H*W-1,2] int32 (increasing edge index; may be NULL when only the mask is wanted);
 *   tree_mask [B,H*W] uint8 (may be NULL): per vertex, bit 0/1/2/3 set when the tree holds the edge to the pixel
 *   above / left / right / below -- the compact form fedicra_bfs_grid consumes.  Any alignment; its bits are set with
 *   32-bit atomics on the aligned words that contain it (up to 3 bytes either side are read and written back unchanged,
 *   so the buffer must not end at the very end of an allocation that is not a multiple of 4 bytes). */
int fedicra_mst_grid(const float *embed, int B, int D, int H, int W, int *edge_out, unsigned char *tree_mask,
                     void *workspace, size_t workspace_bytes, void *stream);

/* The same for the trees of K embeddings of one batch in ONE launch (the image and the decoder heads of the multiscale
 * loss, flower_common.py:784-803: four MinimumSpanningTree calls per step): h_embeds / h_D are HOST arrays of K device
 * pointers [B,D_k,H,W] and channel counts (K <= 8).  Tree k*B + b is image b of embedding k; edge_out [K*B,H*W-1,2],
 * tree_mask [K*B,H*W]; workspace for K*B images. */
int fedicra_mst_grid_multi(const float *const *h_embeds, const int *h_D, int K, int B, int H, int W, int *edge_out,
                           unsigned char *tree_mask, void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------------------------------------------------------------ BFS ---------- */
size_t fedicra_bfs_workspace_bytes(int B, int V);

/* bfs_forward (LTF/src/bfs/bfs.cu:100-135) with max_adj_per_vertex = 4, from root vertex 0.
 *   tree [B,V-1,2]  ->  sorted_index [B,V], sorted_parent [B,V], sorted_child [B,V,4]   (bfs.cu:46-98 semantics)
 *   level_start [B,V+1] first BFS position of every depth (entry n_levels[b] = V), n_levels [B]  (extra outputs)
 * The order is deterministic: by depth, then by parent position, then by vertex id (the reference's is whatever its
 * atomics produce). */
int fedicra_bfs_forward(const int *tree, int B, int V, int *sorted_index, int *sorted_parent, int *sorted_child,
                        int *level_start, int *n_levels, void *workspace, size_t workspace_bytes, void *stream);

/* Same outputs for a tree given as the 4-bit adjacency mask of fedicra_mst_grid (images up to ~440 k pixels; larger
 * ones are refused, use fedicra_bfs_forward).  No workspace.  One warp per image walks the depths out of shared memory. */
int fedicra_bfs_grid(const unsigned char *tree_mask, int B, int H, int W, int *sorted_index, int *sorted_parent,
                     int *sorted_child, int *level_start, int *n_levels, void *stream);

/* Same outputs again, without any per-depth step: Euler tour of the tree ranked by a splitter walk, depth and preorder
 * from one prefix sum along the tour, BFS position = level start + rank of the preorder index inside the level
 * (csrc/euler.cu).  Images up to 262 144 pixels (larger ones: fedicra_bfs_forward); ~100 bytes of workspace per vertex. */
size_t fedicra_bfs_euler_workspace_bytes(int B, int H, int W);
int fedicra_bfs_euler(const unsigned char *tree_mask, int B, int H, int W, int *sorted_index, int *sorted_parent,
                      int *sorted_child, int *level_start, int *n_levels, void *workspace, size_t workspace_bytes,
                      void *stream);
/* The same order started from another root vertex (the reference always starts from vertex 0, bfs.cu:69-74).  The tree
 * filter (refine.cu) does not depend on the root -- out[i] = sum_j S(i,j) f_j / sum_j S(i,j) with S the product of edge
 * weights along the tree path -- but the tree's shape does: rooted at the centre of the image it is 15-40 % shallower
 * than rooted at the corner, and as much wider.  (TreeFilter2D keeps vertex 0: with this library's band sizes the wider
 * levels cost more than the missing ones save; see DESIGN.md.) */
int fedicra_bfs_euler_rooted(const unsigned char *tree_mask, int B, int H, int W, int root, int *sorted_index,
                             int *sorted_parent, int *sorted_child, int *level_start, int *n_levels, void *workspace,
                             size_t workspace_bytes, void *stream);

/* Chain order: a second valid ordering of the same rooted tree (root = vertex 0; bfs.cu:46-98 semantics hold: parents
 * precede children, sorted_parent[0] = 0, child slot value 0 = none), built for fedicra_chain_filter_*: vertices sorted
 * by LIGHT DEPTH (number of non-heaviest-child edges above them), every heavy chain on consecutive positions top to
 * bottom; sorted_child[i] = (heavy child = i + 1 or 0, light children by increasing vertex id, 0...).
 *   group_start [B,33]: first position of every light depth; entries past the last group = V.
 * The reference has one order (its BFS); the filter result does not depend on which valid order is used
 * (refine.cu:19-199 only follow parent / child links).  Same size limit and workspace as fedicra_bfs_euler. */
int fedicra_chain_order(const unsigned char *tree_mask, int B, int H, int W, int *sorted_index, int *sorted_parent,
                        int *sorted_child, int *group_start, void *workspace, size_t workspace_bytes, void *stream);
/* The same for the K*B trees of K embeddings of one batch (h_embeds[k] = [B,D_k,H,W] on the device; tree t belongs to
 * embedding t / B) and, in the same pass, the filter's edge weights of every tree in chain order,
 * edge_weight [K*B,V] = exp(-|e[v(i)] - e[v(par i)]|^2 * h_inv_sigma[k])  (TreeFilter2D.build_edge_weight,
 * LTF/modules/tree_filter.py:91-108; same arithmetic as fedicra_edge_weight_fwd).  h_* are HOST arrays of K entries. */
int fedicra_chain_order_weights(const unsigned char *tree_mask, const float *const *h_embeds, const int *h_D,
                                const float *h_inv_sigma, int K, int B, int H, int W, int *sorted_index,
                                int *sorted_parent, int *sorted_child, int *group_start, float *edge_weight,
                                void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------------------------------------------------------------ tree filter -- */
/* TreeFilter2D.build_edge_weight (LTF/modules/tree_filter.py:91-108): ew[b,i] = exp(-|e[v(i)] - e[v(par i)]|^2 *
 * inv_sigma) in BFS order; embed [B,D,V] in vertex order. */
int fedicra_edge_weight_fwd(const float *embed, const int *sorted_index, const int *sorted_parent, int B, int D, int V,
                            float inv_sigma, float *edge_weight, void *stream);
/* Its backward (what autograd derives from the gathers at tree_filter.py:84-89,97-106): grad_embed [B,D,V] is
 * OVERWRITTEN with the gradient; deterministic (no atomics). sorted_child is needed to visit children. */
int fedicra_edge_weight_bwd(const float *embed, const int *sorted_index, const int *sorted_parent,
                            const int *sorted_child, const float *edge_weight, const float *grad_edge_weight, int B,
                            int D, int V, float inv_sigma, float *grad_embed, void *stream);
/* The same for the trees of K <= 8 embeddings of one batch ([B,D_k,V] each, own orderings and edge weights) in one
 * launch -- the decoder heads of the multiscale loss (flower_common.py:786-806).  The h_* arrays are HOST arrays of K
 * device pointers / values. */
int fedicra_edge_weight_bwd_multi(const float *const *h_embed, const int *const *h_sorted_index,
                                  const int *const *h_sorted_parent, const int *const *h_sorted_child,
                                  const float *const *h_edge_weight, const float *const *h_grad_edge_weight,
                                  const int *h_D, const float *h_inv_sigma, int K, int B, int V,
                                  float *const *h_grad_embed, void *stream);

/* refine_forward (LTF/src/refine/refine.cu:201-249).  Writes 0 into edge_weight[:,0] like the reference (:43-46).
 *   feature_in [B,C,V] vertex order; edge_weight [B,V] BFS order
 *   -> feature_out, feature_aggr [B,C,V] vertex order; feature_aggr_up [B,C,V] BFS order;
 *      weight_sum [B,V] vertex order; weight_sum_up [B,V] BFS order. */
int fedicra_refine_forward(const float *feature_in, float *edge_weight, const int *sorted_index,
                           const int *sorted_parent, const int *sorted_child, const int *level_start,
                           const int *n_levels, int B, int C, int V, float *feature_out, float *feature_aggr,
                           float *feature_aggr_up, float *weight_sum, float *weight_sum_up, void *stream);

size_t fedicra_refine_backward_workspace_bytes(int B, int C, int V);

/* refine_backward_feature (refine.cu:251-300): grad_feature [B,C,V] vertex order. */
int fedicra_refine_backward_feature(const float *edge_weight, const int *sorted_index, const int *sorted_parent,
                                    const int *sorted_child, const int *level_start, const int *n_levels,
                                    const float *weight_sum, const float *grad_out, int B, int C, int V,
                                    float *grad_feature, void *workspace, size_t workspace_bytes, void *stream);

/* refine_backward_weight (refine.cu:302-370): grad_weight [B,V] BFS order. */
int fedicra_refine_backward_weight(const float *edge_weight, const int *sorted_index, const int *sorted_parent,
                                   const int *sorted_child, const int *level_start, const int *n_levels,
                                   const float *feature_out, const float *feature_aggr, const float *feature_aggr_up,
                                   const float *weight_sum, const float *weight_sum_up, const float *grad_out, int B,
                                   int C, int V, float *grad_weight, void *workspace, size_t workspace_bytes,
                                   void *stream);

/* Depth-banded implementation of the same filter (csrc/bands.cu): identical results within fp32 round-off, but the
 * ~1 500-step dependency chain of a level-synchronous sweep is replaced by a few short parallel phases.  This is what
 * TreeFilter2D uses; the three refine_* entry points above keep the reference's tensor-by-tensor interface.
 *   forward : refine_forward (refine.cu:201-249).  feature_aggr / weight_sum (vertex order) may be NULL.  `saved`
 *             (fedicra_tree_filter_saved_bytes) carries what backward needs; edge_weight[:,0] is set to 0.
 *   backward: refine_backward_feature + refine_backward_weight (refine.cu:251-370) in one pass; grad_weight [B,V] (BFS
 *             order) may be NULL (low-level tree: no weight gradient, LTF/functions/refine.py:32-33).
 *   bands   : the band plan of the trees, fedicra_band_plan_ints(B) ints written by fedicra_band_plan from the BFS level
 *             structure (consecutive depths grouped so that a band fits shared memory); one plan per topology. */
size_t fedicra_band_plan_ints(int B);
int fedicra_band_plan(const int *level_start, const int *n_levels, int B, int V, int *bands, void *stream);
size_t fedicra_tree_filter_saved_bytes(int B, int C, int V);
size_t fedicra_tree_filter_workspace_bytes(int B, int C, int V);
int fedicra_tree_filter_forward(const float *feature_in, float *edge_weight, const int *sorted_index,
                                const int *sorted_parent, const int *sorted_child, const int *level_start,
                                const int *n_levels, const int *bands, int B, int C, int V, float *feature_out,
                                float *feature_aggr, float *weight_sum, void *saved, size_t saved_bytes,
                                void *workspace, size_t workspace_bytes, void *stream);
int fedicra_tree_filter_backward(const float *grad_out, const float *edge_weight, const int *sorted_index,
                                 const int *sorted_parent, const int *sorted_child, const int *level_start,
                                 const int *n_levels, const int *bands, int B, int C, int V, const void *saved,
                                 float *grad_feature, float *grad_weight, void *workspace, size_t workspace_bytes,
                                 void *stream);

/* The same filter (refine_forward / refine_backward_feature / refine_backward_weight, LTF/src/refine/refine.cu:201-370)
 * for trees in CHAIN ORDER (fedicra_chain_order): per light depth one scan of affine maps x -> a x + b along the heavy
 * chains, so the sequential dimension is the number of light depths (8-11 at 256^2..512^2) instead of the tree depth
 * (~1 500).  One launch per call (csrc/chains.cu); 1..4 feature channels.
 *   feature_in / grad_out / feature_out / grad_feature [B,C,V] vertex order; edge_weight, grad_weight [B,V] chain order
 *   (edge_weight[b,0] is set to 0 in place like refine.cu:43-46; grad_weight may be NULL);
 *   saved = fedicra_chain_filter_saved_offset bytes of call-private counters, then the leaf->root values and the
 *   root->leaf aggregates, each [B,C+1,V] float in chain order, channel C = the all-ones channel -- i.e.
 *   feature_aggr_up | weight_sum_up and feature_aggr | weight_sum of the reference (refine.cu:201-249);
 *   workspace (backward only): scratch, initialised by the call itself. */
int fedicra_chain_filter_max_channels(void);
size_t fedicra_chain_filter_saved_bytes(int B, int C, int V);
size_t fedicra_chain_filter_saved_offset(int B, int C, int V);
size_t fedicra_chain_filter_workspace_bytes(int B, int C, int V);
int fedicra_chain_filter_forward(const float *feature_in, float *edge_weight, const int *sorted_index,
                                 const int *sorted_parent, const int *sorted_child, const int *group_start, int B, int C,
                                 int V, float *feature_out, void *saved, size_t saved_bytes, void *stream);
int fedicra_chain_filter_backward(const float *grad_out, float *edge_weight, const int *sorted_index,
                                  const int *sorted_parent, const int *sorted_child, const int *group_start, int B, int C,
                                  int V, const void *saved, float *grad_feature, float *grad_weight, void *workspace,
                                  size_t workspace_bytes, void *stream);
/* The same two calls for buffers the caller has ARMED ahead of time: every intermediate word of a chain-filter launch
 * must read "all bits set" when the kernel starts, which the plain calls do with a memset in front of the kernel -- on
 * the critical path of a chain of filters (flower_common.py:786-806 runs four in a row, each consuming the previous
 * one's output).  fedicra_chain_filter_arm(buffers, bytes) sets any number of `saved` / `workspace` buffers at once
 * (one contiguous range), e.g. all eight of a training step before the first filter; each armed buffer serves ONE
 * launch.  (No counterpart in the reference: its kernels keep no cross-CTA state.) */
int fedicra_chain_filter_arm(void *buffers, size_t bytes, void *stream);
int fedicra_chain_filter_forward_armed(const float *feature_in, float *edge_weight, const int *sorted_index,
                                       const int *sorted_parent, const int *sorted_child, const int *group_start, int B,
                                       int C, int V, float *feature_out, void *saved, size_t saved_bytes, void *stream);
int fedicra_chain_filter_backward_armed(const float *grad_out, float *edge_weight, const int *sorted_index,
                                        const int *sorted_parent, const int *sorted_child, const int *group_start, int B,
                                        int C, int V, const void *saved, float *grad_feature, float *grad_weight,
                                        void *workspace, size_t workspace_bytes, void *stream);

/* F.interpolate(x, size=(H, W), mode='bilinear', align_corners=False) of up to 4 decoder heads in one launch
 * (flower_common.py:789,796,803) and its gradient.  h_src[k] = [planes, h_k, w_k], h_dst[k] = [planes, H, W] (forward);
 * h_grad_out[k] = [planes, H, W], h_grad_in[k] = [planes, h_k, w_k] (backward); the h_* arrays are HOST arrays.  The
 * forward reproduces torch's CUDA result bit for bit (the MST is built on these values); the backward gathers in a fixed
 * order where torch scatters with atomics. */
int fedicra_bilinear_resize_fwd(const float *const *h_src, float *const *h_dst, const int *h_h, const int *h_w, int n,
                                int planes, int H, int W, void *stream);
int fedicra_bilinear_resize_bwd(const float *const *h_grad_out, float *const *h_grad_in, const int *h_h, const int *h_w,
                                int n, int planes, int H, int W, void *stream);

/* ------------------------------------------------------------------------------------------------ loss glue ---- */
/* prob = softmax over the channel axis of x [B,C,P]  (flower_common.py:781, torch.softmax(preds, dim=1)); C <= 16. */
int fedicra_channel_softmax_fwd(const float *x, int B, int C, long long P, float *y, void *stream);
/* Its backward: grad_x = y * (grad_y - sum_c grad_y * y). */
int fedicra_channel_softmax_bwd(const float *y, const float *grad_y, int B, int C, long long P, float *grad_x,
                                void *stream);

/* Masked L1 tail of the tree-energy losses (flower_common.py:777, 813-818) in one pass:
 *   sums[0] = sum_{b,c,p} roi[b,p] * |prob[b,c,p] - filtered[b,c,p]|,   sums[1] = N = sum roi,
 *   loss[0] = weight * sums[0] / max(N, 1)      (equals the reference's `if N > 0: loss /= N` without its host sync).
 * roi [B,P] bytes (0 / non-0).  The workspace (fedicra_masked_l1_workspace_bytes) must be zero before its first use;
 * the call leaves it reusable.  Deterministic (fixed summation order). */
size_t fedicra_masked_l1_workspace_bytes(void);
int fedicra_masked_l1_fwd(const float *prob, const float *filtered, const unsigned char *roi, int B, int C, long long P,
                          float weight, float *sums, float *loss, void *workspace, size_t workspace_bytes, void *stream);
/* Backward: grad_prob = s * roi * sign(prob - filtered), grad_filtered = -grad_prob (either may be NULL), with
 * s = grad_out[0] * weight / max(n[0], 1) when n is given (gradient of loss[0]), s = grad_out[0] when n is NULL
 * (gradient of sums[0]; the sharded path divides after its all-reduce). */
int fedicra_masked_l1_bwd(const float *prob, const float *filtered, const unsigned char *roi, const float *grad_out,
                          const float *n, float weight, int B, int C, long long P, float *grad_prob,
                          float *grad_filtered, void *stream);

/* Telemetry of the resident filter kernel (no counterpart in the reference): with FEDICRA_PROF=1 in the environment the
 * CTAs of the first image of a launch record %globaltimer at their phase boundaries.  Copies the last launch's
 * [bands + 1][12] nanosecond stamps to the HOST buffer h_out (n = its capacity in elements; the required count is
 * returned; synchronises the device). */
int fedicra_debug_filter_profile(long long *h_out, int n);
int fedicra_debug_chain_profile(long long *h_out, int n);   /* same for csrc/chains.cu: returns the value count */
int fedicra_debug_mst_profile(long long *h_out, int n);     /* clock64 stamps of tree 0's MST contraction (csrc/mst.cu) */

/* ------------------------------------------------------------------------------------------------ aggregation -- */
/* Adaptive local aggregation, element-wise part of MyModel.set_weights (flower_common.py:590-602) over one flat
 * array: w <- clamp(w - eta*g*(s-p), 0, 1);  t <- p + (s-p)*w.   No fused multiply-add: bit-exact with torch.
 * g may be NULL (blend only, flower_common.py:553-555). */
int fedicra_ala_update_flat(const float *p, const float *s, const float *g, float *w, float *t, int64_t n, float eta,
                            void *stream);

/* Same over a list of tensors in ONE launch.  Tables are device arrays: p/s/g/w/t [n_tensors] of pointers (g entries
 * may be 0 = "no gradient, skip the weight update", flower_common.py:593-594), chunk_tensor/chunk_begin [n_chunks]
 * int32/int64 describing which tensor and which element offset each CTA works on, sizes [n_tensors] int64. */
int fedicra_ala_update_multi(const float *const *p, const float *const *s, const float *const *g, float *const *w,
                             float *const *t, const int64_t *sizes, const int *chunk_tensor,
                             const int64_t *chunk_begin, int n_chunks, int chunk_elems, float eta, void *stream);

/* FedAvg pre-scale: x <- x * alpha (alpha = n_i / sum n), the local half of sum_i n_i W_i / sum_i n_i
 * (flwr 1.0.0 aggregate(), called from flower_common.py:262); the sum itself is one NCCL all-reduce. */
int fedicra_scale_inplace(float *x, int64_t n, float alpha, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* FEDICRA_SM100_H */
