// GPU ball-tree build that reproduces the reference permutation bit for bit.
//
// Reference: models/components/balltree/balltree.h:32-66 (find_split_dimension), :82-101
// (partition_node_indices == std::nth_element), :121-186 (build_tree), :205-259
// (partition_ball_tree); balltree.pyx:58-60 (depth), :576-596 (rotation matrix), :598-662
// (build_balltree_with_rotations).  The element order std::nth_element leaves behind is part of
// the observable output, so GCC-13 libstdc++'s introselect (bits/stl_algo.h: __introselect,
// __unguarded_partition_pivot, __move_median_to_first, __heap_select, __insertion_sort) is
// restated here step for step on a shared-memory index array.
//
// Mapping: one CTA per tree; the tree is built level-synchronously (all nodes of a level in
// parallel, one thread per node; the node ranges follow from mid = (start+end)/2 alone).
// Keys are the fp32 positions: the reference promotes them to fp64 before comparing
// (balltree.pyx:542), which is exact and monotonic, so fp32 compares give the same order; the
// max-spread test (balltree.h:52-62) subtracts in fp64 and is done in fp64 here.
#include <stdlib.h>

#include "common.cuh"

namespace haet {

struct KeyCtx {
    const float *key;  // key of element i: key[i * d + dim]
    int d, dim;
};
__device__ __forceinline__ float KEY(const KeyCtx &c, int i) { return c.key[i * c.d + c.dim]; }
__device__ __forceinline__ bool LESS(const KeyCtx &c, int a, int b) { return KEY(c, a) < KEY(c, b); }
__device__ __forceinline__ void SWAPI(int *a, int *b) { int t = *a; *a = *b; *b = t; }

// stl_algo.h __move_median_to_first
__device__ void move_median_to_first(const KeyCtx &c, int *result, int *a, int *b, int *cc) {
    if (LESS(c, *a, *b)) {
        if (LESS(c, *b, *cc)) SWAPI(result, b);
        else if (LESS(c, *a, *cc)) SWAPI(result, cc);
        else SWAPI(result, a);
    } else if (LESS(c, *a, *cc)) SWAPI(result, a);
    else if (LESS(c, *b, *cc)) SWAPI(result, cc);
    else SWAPI(result, b);
}

// stl_algo.h __unguarded_partition
__device__ int *unguarded_partition(const KeyCtx &c, int *first, int *last, int *pivot) {
    const float pv = KEY(c, *pivot);
    for (;;) {
        while (KEY(c, *first) < pv) ++first;
        --last;
        while (pv < KEY(c, *last)) --last;
        if (!(first < last)) return first;
        SWAPI(first, last);
        ++first;
    }
}

// stl_heap.h __push_heap / __adjust_heap / __make_heap, stl_algo.h __heap_select
__device__ void push_heap_(const KeyCtx &c, int *first, int hole, int top, int value) {
    int parent = (hole - 1) / 2;
    while (hole > top && LESS(c, first[parent], value)) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}
__device__ void adjust_heap_(const KeyCtx &c, int *first, int hole, int len, int value) {
    const int top = hole;
    int second = hole;
    while (second < (len - 1) / 2) {
        second = 2 * (second + 1);
        if (LESS(c, first[second], first[second - 1])) second--;
        first[hole] = first[second];
        hole = second;
    }
    if ((len & 1) == 0 && second == (len - 2) / 2) {
        second = 2 * (second + 1);
        first[hole] = first[second - 1];
        hole = second - 1;
    }
    push_heap_(c, first, hole, top, value);
}
__device__ void heap_select_(const KeyCtx &c, int *first, int *middle, int *last) {
    int len = (int)(middle - first);
    if (len >= 2) {
        int parent = (len - 2) / 2;
        for (;;) {
            adjust_heap_(c, first, parent, len, first[parent]);
            if (parent == 0) break;
            parent--;
        }
    }
    for (int *i = middle; i < last; ++i) {
        if (LESS(c, *i, *first)) {
            int value = *i;
            *i = *first;
            adjust_heap_(c, first, 0, len, value);
        }
    }
}

// stl_algo.h __insertion_sort (+ __unguarded_linear_insert)
__device__ void insertion_sort_(const KeyCtx &c, int *first, int *last) {
    if (first == last) return;
    for (int *i = first + 1; i != last; ++i) {
        int val = *i;
        if (LESS(c, val, *first)) {
            for (int *p = i; p != first; --p) *p = *(p - 1);
            *first = val;
        } else {
            int *lastp = i, *next = i - 1;
            while (LESS(c, val, *next)) {
                *lastp = *next;
                lastp = next;
                --next;
            }
            *lastp = val;
        }
    }
}

// std::nth_element == __introselect(first, nth, last, 2 * __lg(last - first))
__device__ void nth_element_(const KeyCtx &c, int *first, int *nth, int *last) {
    if (first == last || nth == last) return;
    int depth_limit = 2 * (31 - __clz((int)(last - first)));
    while (last - first > 3) {
        if (depth_limit == 0) {
            heap_select_(c, first, nth + 1, last);
            SWAPI(first, nth);
            return;
        }
        --depth_limit;
        int *mid = first + (last - first) / 2;
        move_median_to_first(c, first, first + 1, mid, last - 1);
        int *cut = unguarded_partition(c, first + 1, last, first);
        if (cut <= nth) first = cut;
        else last = cut;
    }
    insertion_sort_(c, first, last);
}

// balltree.h:32-66 -- strict '>' keeps the lowest dimension on ties; spreads in fp64
__device__ int find_split_dimension(const float *key, const int *idx, int s, int e, int d) {
    if (e - s <= 1) return 0;
    int best = 0;
    double max_spread = 0;
    for (int dim = 0; dim < d; ++dim) {
        float mn = key[idx[s] * d + dim], mx = mn;
        for (int i = s + 1; i < e; ++i) {
            float v = key[idx[i] * d + dim];
            mn = v < mn ? v : mn;
            mx = mx < v ? v : mx;
        }
        double spread = (double)mx - (double)mn;
        if (spread > max_spread) { max_spread = spread; best = dim; }
    }
    return best;
}

__device__ __forceinline__ void node_range(int j, int level, int n, int &s, int &e) {
    s = 0; e = n;
    for (int b = level - 1; b >= 0; --b) {
        int mid = (s + e) / 2;
        if ((j >> b) & 1) s = mid; else e = mid;
    }
}

// ---- warp-cooperative versions for the large nodes near the root ------------------------------------------
// find_split_dimension: min / max are order independent, so a strided scan + shuffle reduction is exact.
__device__ int warp_find_split_dimension(const float *key, const int *idx, int s, int e, int d) {
    if (e - s <= 1) return 0;
    const int lane = threadIdx.x & 31;
    int best = 0;
    double max_spread = 0;
    for (int dim = 0; dim < d; ++dim) {
        float mn = INFINITY, mx = -INFINITY;
        for (int i = s + lane; i < e; i += 32) {
            const float v = key[idx[i] * d + dim];
            mn = v < mn ? v : mn;
            mx = mx < v ? v : mx;
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            const float a = __shfl_xor_sync(0xffffffffu, mn, o), b = __shfl_xor_sync(0xffffffffu, mx, o);
            mn = a < mn ? a : mn;
            mx = mx < b ? b : mx;
        }
        const double spread = (double)mx - (double)mn;
        if (spread > max_spread) { max_spread = spread; best = dim; }
    }
    return best;
}

// std::nth_element with the Hoare partition of every introselect round done by the whole warp.
// __unguarded_partition(f, last, pivot) is a deterministic function of the range: with A = positions (ascending,
// from f) whose key is >= pivot and B = positions (descending, from last - 1) whose key is <= pivot, the
// sequential loop swaps A[k] <-> B[k] for k < K, K = #{k : A[k] < B[k]} (a swapped position is never visited
// again), and returns min(A[K], B[K-1]).  Ranks come from ballots; the result is bit-identical to the sequential
// code (the median-of-3, the heap-select fallback and the final insertion sort stay sequential on lane 0).
__device__ void warp_nth_element(const KeyCtx &c, int *idx, int first, int nth, int last, int *Apos, int *Bpos) {
    const int lane = threadIdx.x & 31;
    if (first == last || nth == last) return;
    int depth_limit = 2 * (31 - __clz(last - first));
    while (last - first > 3) {
        if (depth_limit == 0) {
            if (lane == 0) {
                heap_select_(c, idx + first, idx + nth + 1, idx + last);
                SWAPI(idx + first, idx + nth);
            }
            __syncwarp();
            return;
        }
        --depth_limit;
        if (lane == 0) move_median_to_first(c, idx + first, idx + first + 1, idx + first + (last - first) / 2, idx + last - 1);
        __syncwarp();
        const float pv = KEY(c, idx[first]);
        const int f = first + 1;
        int cntA = 0, cntB = 0;
        for (int base = f; base < last; base += 32) {
            const int p = base + lane;
            const bool in = p < last;
            const float k = in ? KEY(c, idx[p]) : 0.f;
            const bool isA = in && !(k < pv), isB = in && !(pv < k);
            const unsigned mA = __ballot_sync(0xffffffffu, isA), mB = __ballot_sync(0xffffffffu, isB);
            const unsigned lt = (1u << lane) - 1u;
            if (isA) Apos[first + cntA + __popc(mA & lt)] = p;
            if (isB) Bpos[first + cntB + __popc(mB & lt)] = p;   // ascending; B[k] (descending) = Bpos[first + cntB - 1 - k]
            cntA += __popc(mA);
            cntB += __popc(mB);
        }
        __syncwarp();
        const int lim = cntA < cntB ? cntA : cntB;
        int K = 0;
        for (int base = 0; base < lim; base += 32) {
            const int k = base + lane;
            const bool ok = k < lim && Apos[first + k] < Bpos[first + cntB - 1 - k];
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            K += __popc(m);
            if (m != 0xffffffffu) break;   // A ascending, B descending: the predicate is monotone in k
        }
        const int aK = K < cntA ? Apos[first + K] : 0x7fffffff;
        const int bK1 = K >= 1 ? Bpos[first + cntB - K] : 0x7fffffff;
        const int cut = aK < bK1 ? aK : bK1;
        for (int k = lane; k < K; k += 32) SWAPI(idx + Apos[first + k], idx + Bpos[first + cntB - 1 - k]);
        __syncwarp();
        if (cut <= nth) first = cut;
        else last = cut;
    }
    if (lane == 0) insertion_sort_(c, idx + first, idx + last);
    __syncwarp();
}

constexpr int BT_WARP_NODE = 12;   // nodes of at least this many points are split by a whole warp

// Split every node of levels [0, n_levels) of the implicit median tree over idx[0, n).
// stop_small: partition_ball_tree stops at <= 1 point (balltree.h:219); build_tree always splits.
__device__ void split_levels(const float *key, int *idx, int n, int d, int n_levels, int *Apos, int *Bpos) {
    const int warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    for (int level = 0; level < n_levels; ++level) {
        const int nodes = 1 << level;
        if ((n >> level) >= BT_WARP_NODE) {
            for (int j = warp; j < nodes; j += n_warps) {
                int s, e;
                node_range(j, level, n, s, e);
                if (e - s > 1) {
                    KeyCtx c{key, d, warp_find_split_dimension(key, idx, s, e, d)};
                    warp_nth_element(c, idx, s, (s + e) / 2, e, Apos, Bpos);
                }
            }
        } else {
            for (int j = threadIdx.x; j < nodes; j += blockDim.x) {
                int s, e;
                node_range(j, level, n, s, e);
                if (e - s > 1) {
                    KeyCtx c{key, d, find_split_dimension(key, idx, s, e, d)};
                    nth_element_(c, idx + s, idx + (s + e) / 2, idx + e);
                }
            }
        }
        __syncthreads();
    }
}

struct BallTreeArgs {
    const float *pos;
    int64_t pts, slots;
    int d, max_level, n_levels, rotate;
    int strides[HAET_MAX_LEVELS];
    int targets[HAET_MAX_LEVELS];
    float R[9];
    int64_t *tree_idx;
    uint8_t *tree_mask;
    int64_t *rot_idx[HAET_MAX_LEVELS];
};

__global__ void __launch_bounds__(256) balltree_kernel(BallTreeArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int n = (int)a.pts, slots = (int)a.slots, d = a.d;
    float *P = reinterpret_cast<float *>(smem_raw);      // [max(n, slots) * d]
    float *rotA = P + (size_t)slots * d;                  // [slots * d]
    float *rotB = rotA + (size_t)slots * d;               // [slots * d]
    int *idx = reinterpret_cast<int *>(rotB + (size_t)slots * d);  // [slots]
    int *leaf = idx + slots;                              // [slots]
    int *Apos = leaf + slots, *Bpos = Apos + slots;       // [slots] each: scratch of the warp-parallel partition
    const int tid = threadIdx.x;
    const int64_t tree = blockIdx.x;
    const int64_t row0 = tree * a.pts;
    // blockIdx.y = role: the rotated partitions of the levels are independent of each other (each starts from the pooled
    // rotated leaves, not from the previous level's permutation), so CTA (tree, r) rebuilds the (deterministic) main tree
    // and then partitions only the levels lvl % gridDim.y == r; role 0 also writes the main tree.  0.111 -> see DESIGN
    const int role = blockIdx.y, roles = gridDim.y;

    for (int i = tid; i < n * d; i += blockDim.x) P[i] = a.pos[row0 * d + i];
    for (int i = tid; i < n; i += blockDim.x) idx[i] = i;
    __syncthreads();

    // ---- main tree (balltree.h:121-186) ----
    split_levels(P, idx, n, d, a.max_level, Apos, Bpos);
    for (int j = tid; j < (1 << a.max_level); j += blockDim.x) {
        int s, e;
        node_range(j, a.max_level, n, s, e);
        const bool pair = (e - s) >= 2;
        const int i0 = idx[s], i1 = pair ? idx[s + 1] : idx[s];
        leaf[2 * j] = i0; leaf[2 * j + 1] = i1;
        if (role == 0) {
            const int64_t o = tree * a.slots + 2 * j;
            a.tree_idx[o] = row0 + i0; a.tree_idx[o + 1] = row0 + i1;
            a.tree_mask[o] = 1; a.tree_mask[o + 1] = pair ? 1 : 0;
        }
    }
    __syncthreads();
    if (!a.rotate) return;

    // ---- rotated partitions (balltree.pyx:639-660) ----
    // leaves @ R in fp32: forward FMA chain, bit-identical to the torch CPU matmul the oracle pins
    for (int i = tid; i < slots * d; i += blockDim.x) {
        const int r = i / d, j = i - r * d;
        const float *lv = P + leaf[r] * d;
        float acc = __fmul_rn(lv[0], a.R[j]);
        for (int k = 1; k < d; ++k) acc = __fmaf_rn(lv[k], a.R[k * d + j], acc);
        rotA[i] = acc;
    }
    __syncthreads();
    int cur = slots;
    for (int lvl = 0; lvl < a.n_levels; ++lvl) {
        if (lvl % roles == role) {   // CTA-uniform
            for (int i = tid; i < cur; i += blockDim.x) idx[i] = i;
            __syncthreads();
            // number of levels that can still split anything: ranges shrink to <= 1 after ceil(log2 cur)
            int lv = a.targets[lvl];
            int maxlv = cur > 1 ? 32 - __clz(cur - 1) : 0;
            if (lv > maxlv) lv = maxlv;
            split_levels(rotA, idx, cur, d, lv, Apos, Bpos);
            int64_t *out = a.rot_idx[lvl] + tree * cur;
            for (int i = tid; i < cur; i += blockDim.x) out[i] = tree * cur + idx[i];
        }
        if (lvl + 1 >= a.n_levels) break;
        bool later = false;   // does this CTA still own a level after lvl?
        for (int l2 = lvl + 1; l2 < a.n_levels; ++l2) later |= (l2 % roles == role);
        if (!later) break;
        if (lvl < a.n_levels - 1) {
            const int s = a.strides[lvl];
            const int nn = cur / s;
            for (int i = tid; i < nn * d; i += blockDim.x) {
                const int r = i / d, j = i - r * d;
                float acc = rotA[(r * s) * d + j];
                for (int k = 1; k < s; ++k) acc = __fadd_rn(acc, rotA[(r * s + k) * d + j]);
                rotB[i] = __fdiv_rn(acc, (float)s);
            }
            __syncthreads();
            float *t = rotA; rotA = rotB; rotB = t;
            cur = nn;
        }
        __syncthreads();
    }
}

static int tree_depth(int64_t n) { return (int)ceil(log2((double)n)) - 1; }  // balltree.pyx:58-60

}  // namespace haet

extern "C" int64_t haet_balltree_slots(int64_t n_trees, int64_t pts) {
    if (pts < 2 || n_trees < 0) return -1;
    return n_trees * (((int64_t)1 << haet::tree_depth(pts)) * 2);
}

extern "C" int haet_balltree_build(const float *pos_d, int64_t n_trees, int64_t pts, int32_t d,
                                   const int32_t *strides, const int32_t *ball_sizes, int32_t n_levels,
                                   double angle_deg, int64_t *tree_idx_d, uint8_t *tree_mask_d,
                                   int64_t *const *rot_idx_d, void *stream) {
    using namespace haet;
    HAET_CHECK_ARG(pts >= 2, "every tree needs at least 2 points (got %lld)", (long long)pts);
    HAET_CHECK_ARG(d >= 1 && d <= 3, "dimensionality must be 1..3 (got %d)", d);
    HAET_CHECK_ARG(n_levels >= 1 && n_levels <= HAET_MAX_LEVELS, "n_levels out of range");
    HAET_CHECK_ARG(pts <= 8192, "trees of more than 8192 points are not supported");
    if (n_trees == 0) return HAET_OK;
    BallTreeArgs a;
    a.pos = pos_d; a.pts = pts; a.d = d; a.n_levels = n_levels;
    a.max_level = tree_depth(pts);
    a.slots = ((int64_t)1 << a.max_level) * 2;
    a.rotate = angle_deg > 0 ? 1 : 0;
    a.tree_idx = tree_idx_d; a.tree_mask = tree_mask_d;
    const int64_t total = n_trees * a.slots;
    int64_t cur = a.slots;
    for (int i = 0; i < n_levels; ++i) {
        // balltree.pyx:643: max(0, int(math.log2(leaves.shape[0] / bs))) with the TOTAL leaf count
        HAET_CHECK_ARG(ball_sizes[i] > 0, "ball sizes must be positive");
        int t = (int)log2((double)total / (double)ball_sizes[i]);
        a.targets[i] = t < 0 ? 0 : t;
        a.strides[i] = 1;
        a.rot_idx[i] = (a.rotate && rot_idx_d) ? rot_idx_d[i] : nullptr;
        if (a.rotate) HAET_CHECK_ARG(a.rot_idx[i] != nullptr, "rot_idx_d[%d] is null", i);
        if (i < n_levels - 1) {
            HAET_CHECK_ARG(strides[i] >= 1 && cur % strides[i] == 0, "stride %d does not divide %lld leaves",
                           strides[i], (long long)cur);
            a.strides[i] = strides[i];
            cur /= strides[i];
        }
    }
    // generate_rotation_matrix (balltree.pyx:576-596): python double math stored as fp32
    const double ang = angle_deg * (M_PI / 180.0), c = cos(ang), s = sin(ang);
    for (int i = 0; i < 9; ++i) a.R[i] = 0.f;
    if (d == 1) a.R[0] = 1.f;  // the reference has no 1-D matrix; identity keeps the call defined
    else if (d == 2) { a.R[0] = (float)c; a.R[1] = (float)(-s); a.R[2] = (float)s; a.R[3] = (float)c; }
    else {
        a.R[0] = (float)(c * c); a.R[1] = (float)(s * c * (s - 1)); a.R[2] = (float)(s * (s + c * c));
        a.R[3] = (float)(s * c); a.R[4] = (float)(s * s * s + c * c); a.R[5] = (float)(s * c * (s - 1));
        a.R[6] = (float)(-s); a.R[7] = (float)(s * c); a.R[8] = (float)(c * c);
    }
    const size_t smem = (size_t)a.slots * d * 3 * sizeof(float) + (size_t)a.slots * 4 * sizeof(int);
    if (smem > 48 * 1024) HAET_DYN_SMEM(balltree_kernel, smem);
    int roles = a.rotate ? n_levels : 1;   // one CTA per (tree, rotated level)
    if (roles > 4) roles = 4;
    if (options().tree_roles > 0) { roles = options().tree_roles; if (roles > n_levels) roles = 1; }
    balltree_kernel<<<dim3((unsigned)n_trees, (unsigned)roles), 256, smem, (cudaStream_t)stream>>>(a);
    HAET_LAUNCHED();
    return HAET_OK;
}
