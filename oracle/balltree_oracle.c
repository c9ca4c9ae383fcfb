/*
 * balltree_oracle.c -- TEST INFRASTRUCTURE ONLY (the checker, never the product).
 *
 * Plain-C CPU restatement of the reference ball-tree construction used by
 * Erwin inside HAET.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.
 *
 * What it follows (all paths relative to /root/reference):
 *   models/components/balltree/balltree.h:32-66     find_split_dimension
 *   models/components/balltree/balltree.h:82-101    partition_node_indices (std::nth_element)
 *   models/components/balltree/balltree.h:121-186   build_tree
 *   models/components/balltree/balltree.h:205-259   partition_ball_tree
 *   models/components/balltree/balltree.pyx:58-60   compute_tree_depth
 *   models/components/balltree/balltree.pyx:89-122  build_batch_trees (+ batch.start offset)
 *   models/components/balltree/balltree.pyx:154-183 partition_batch_trees
 *   models/components/balltree/balltree.pyx:576-596 generate_rotation_matrix
 *   models/components/balltree/balltree.pyx:598-662 build_balltree_with_rotations
 * and, because the reference delegates the partition to std::nth_element, the
 * published algorithm of GCC 13 libstdc++ (third-party, not under /root/reference):
 *   bits/stl_algo.h  __move_median_to_first, __unguarded_partition,
 *                    __unguarded_partition_pivot, __introselect, __heap_select,
 *                    __insertion_sort, __unguarded_linear_insert
 *   bits/stl_heap.h  __push_heap, __adjust_heap, __pop_heap, __make_heap
 * The element order std::nth_element leaves behind is part of the reference's
 * observable output, so the introselect is restated step for step.
 *
 * Parity status: PINNED -- checked bit-for-bit against the reference's own
 * compiled extension (oracle/_ref/balltree*.so, built by oracle/build_ref.py)
 * in tests/test_oracle_vs_reference.py and against tests/golden/balltree_*.npz.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef int64_t idx_t;

/* ---- comparator context: key of point i is data[i*nf + dim] (double) ---- */
typedef struct {
    const double *data;
    idx_t nf;
    idx_t dim;
} keyctx;

static inline double KEY(const keyctx *c, idx_t i) { return c->data[i * c->nf + c->dim]; }
static inline int LESS(const keyctx *c, idx_t a, idx_t b) { return KEY(c, a) < KEY(c, b); }
static inline void SWAP(idx_t *a, idx_t *b) { idx_t t = *a; *a = *b; *b = t; }

/* stl_algo.h __move_median_to_first */
static void move_median_to_first(const keyctx *c, idx_t *result, idx_t *a, idx_t *b, idx_t *cc)
{
    if (LESS(c, *a, *b)) {
        if (LESS(c, *b, *cc)) SWAP(result, b);
        else if (LESS(c, *a, *cc)) SWAP(result, cc);
        else SWAP(result, a);
    } else if (LESS(c, *a, *cc)) SWAP(result, a);
    else if (LESS(c, *b, *cc)) SWAP(result, cc);
    else SWAP(result, b);
}

/* stl_algo.h __unguarded_partition */
static idx_t *unguarded_partition(const keyctx *c, idx_t *first, idx_t *last, idx_t *pivot)
{
    for (;;) {
        while (LESS(c, *first, *pivot)) ++first;
        --last;
        while (LESS(c, *pivot, *last)) --last;
        if (!(first < last)) return first;
        SWAP(first, last);
        ++first;
    }
}

/* stl_algo.h __unguarded_partition_pivot */
static idx_t *unguarded_partition_pivot(const keyctx *c, idx_t *first, idx_t *last)
{
    idx_t *mid = first + (last - first) / 2;
    move_median_to_first(c, first, first + 1, mid, last - 1);
    return unguarded_partition(c, first + 1, last, first);
}

/* stl_heap.h __push_heap (comparator: iter < value) */
static void push_heap_(const keyctx *c, idx_t *first, idx_t hole, idx_t top, idx_t value)
{
    idx_t parent = (hole - 1) / 2;
    while (hole > top && LESS(c, first[parent], value)) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}

/* stl_heap.h __adjust_heap */
static void adjust_heap_(const keyctx *c, idx_t *first, idx_t hole, idx_t len, idx_t value)
{
    const idx_t top = hole;
    idx_t second = hole;
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

/* stl_heap.h __make_heap */
static void make_heap_(const keyctx *c, idx_t *first, idx_t *last)
{
    idx_t len = last - first;
    if (len < 2) return;
    idx_t parent = (len - 2) / 2;
    for (;;) {
        idx_t value = first[parent];
        adjust_heap_(c, first, parent, len, value);
        if (parent == 0) return;
        parent--;
    }
}

/* stl_algo.h __heap_select (with stl_heap.h __pop_heap inlined) */
static void heap_select_(const keyctx *c, idx_t *first, idx_t *middle, idx_t *last)
{
    make_heap_(c, first, middle);
    for (idx_t *i = middle; i < last; ++i) {
        if (LESS(c, *i, *first)) {
            idx_t value = *i;
            *i = *first;
            adjust_heap_(c, first, 0, middle - first, value);
        }
    }
}

/* stl_algo.h __insertion_sort + __unguarded_linear_insert */
static void insertion_sort_(const keyctx *c, idx_t *first, idx_t *last)
{
    if (first == last) return;
    for (idx_t *i = first + 1; i != last; ++i) {
        if (LESS(c, *i, *first)) {
            idx_t val = *i;
            memmove(first + 1, first, (size_t)(i - first) * sizeof(idx_t));
            *first = val;
        } else {
            idx_t val = *i;
            idx_t *lastp = i;
            idx_t *next = i - 1;
            while (LESS(c, val, *next)) {
                *lastp = *next;
                lastp = next;
                --next;
            }
            *lastp = val;
        }
    }
}

static int lg_(idx_t n) { int k = 0; while (n > 1) { n >>= 1; ++k; } return k; }

/* std::nth_element == __introselect(first, nth, last, 2*__lg(n)) */
static void nth_element_(const keyctx *c, idx_t *first, idx_t *nth, idx_t *last)
{
    if (first == last || nth == last) return;
    idx_t depth_limit = 2 * lg_(last - first);
    while (last - first > 3) {
        if (depth_limit == 0) {
            heap_select_(c, first, nth + 1, last);
            SWAP(first, nth);
            return;
        }
        --depth_limit;
        idx_t *cut = unguarded_partition_pivot(c, first, last);
        if (cut <= nth) first = cut;
        else last = cut;
    }
    insertion_sort_(c, first, last);
}

/* balltree.h:32-66 */
static idx_t find_split_dimension(const double *data, const idx_t *idx, idx_t start, idx_t end, idx_t nf)
{
    if (end - start <= 1) return 0;
    idx_t best = 0;
    double max_spread = 0;
    for (idx_t dim = 0; dim < nf; ++dim) {
        double mn = data[idx[start] * nf + dim], mx = mn;
        for (idx_t i = start + 1; i < end; ++i) {
            double v = data[idx[i] * nf + dim];
            mn = v < mn ? v : mn; /* std::min(min_val, val) */
            mx = mx < v ? v : mx; /* std::max(max_val, val) */
        }
        double spread = mx - mn;
        if (spread > max_spread) { max_spread = spread; best = dim; }
    }
    return best;
}

/* balltree.h:121-186 */
static void build_tree(const double *data, idx_t *idx, idx_t *out, uint8_t *mask,
                       idx_t s, idx_t e, idx_t out_start, idx_t nf, idx_t level, idx_t max_level)
{
    idx_t n = e - s;
    if (level == max_level) {
        if (n == 1) {
            out[out_start] = idx[s]; out[out_start + 1] = idx[s];
            mask[out_start] = 1; mask[out_start + 1] = 0;
        } else {
            out[out_start] = idx[s]; out[out_start + 1] = idx[s + 1];
            mask[out_start] = 1; mask[out_start + 1] = 1;
        }
        return;
    }
    idx_t dim = find_split_dimension(data, idx, s, e, nf);
    idx_t mid = (s + e) / 2;
    if (n > 1) {
        keyctx c = {data, nf, dim};
        nth_element_(&c, idx + s, idx + mid, idx + e);
    }
    idx_t left_size = ((idx_t)1 << (max_level - level - 1)) * 2;
    build_tree(data, idx, out, mask, s, mid, out_start, nf, level + 1, max_level);
    build_tree(data, idx, out, mask, mid, e, out_start + left_size, nf, level + 1, max_level);
}

/* balltree.h:205-259 */
static void partition_tree(const double *data, idx_t *idx, idx_t *out, idx_t s, idx_t e,
                           idx_t out_start, idx_t nf, idx_t level, idx_t target)
{
    idx_t n = e - s;
    if (target == level || n <= 1) {
        for (idx_t i = 0; i < n; ++i) out[out_start + i] = idx[s + i];
        return;
    }
    idx_t dim = find_split_dimension(data, idx, s, e, nf);
    idx_t mid = (s + e) / 2;
    if (n > 1) {
        keyctx c = {data, nf, dim};
        nth_element_(&c, idx + s, idx + mid, idx + e);
    }
    partition_tree(data, idx, out, s, mid, out_start, nf, level + 1, target);
    partition_tree(data, idx, out, mid, e, out_start + (mid - s), nf, level + 1, target);
}

/* balltree.pyx:58-60 */
static idx_t tree_depth(idx_t n) { return (idx_t)ceil(log2((double)n)) - 1; }

/* Number of output slots of the main tree for the given offsets (pyx:256-274). */
int64_t haet_oracle_tree_slots(const int64_t *offsets, int64_t n_trees)
{
    int64_t total = 0;
    for (int64_t t = 0; t < n_trees; ++t) {
        idx_t n = offsets[t + 1] - offsets[t];
        if (n < 2) return -1; /* depth -1: the reference shifts by a negative count (UB) */
        total += ((idx_t)1 << tree_depth(n)) * 2;
    }
    return total;
}

/* build_balltree (pyx:525-550 via :424-472, :185-300, :62-122); data is fp32 promoted to double (pyx:542). */
int haet_oracle_build_balltree(const float *data32, int64_t n_points, int d,
                               const int64_t *offsets, int64_t n_trees,
                               int64_t *tree_idx, uint8_t *tree_mask)
{
    double *data = (double *)malloc((size_t)n_points * d * sizeof(double));
    if (!data) return -1;
    for (int64_t i = 0; i < n_points * d; ++i) data[i] = (double)data32[i];
    idx_t tree_off = 0;
    for (int64_t t = 0; t < n_trees; ++t) {
        idx_t s = offsets[t], e = offsets[t + 1], n = e - s;
        if (n < 2) { free(data); return -2; }
        idx_t max_level = tree_depth(n);
        idx_t slots = ((idx_t)1 << max_level) * 2;
        idx_t *idx = (idx_t *)malloc((size_t)n * sizeof(idx_t));
        for (idx_t i = 0; i < n; ++i) idx[i] = i;
        build_tree(data + s * d, idx, tree_idx + tree_off, tree_mask + tree_off, 0, n, 0, d, 0, max_level);
        for (idx_t i = 0; i < slots; ++i) tree_idx[tree_off + i] += s;
        free(idx);
        tree_off += slots;
    }
    free(data);
    return 0;
}

/* partition_balltree (pyx:552-574 via :474-523, :302-422, :124-183). */
int haet_oracle_partition_balltree(const float *data32, int64_t n_points, int d,
                                   const int64_t *offsets, int64_t n_trees,
                                   int64_t target_level, int64_t *out)
{
    if (target_level < 0) return -3;
    double *data = (double *)malloc((size_t)n_points * d * sizeof(double));
    if (!data) return -1;
    for (int64_t i = 0; i < n_points * d; ++i) data[i] = (double)data32[i];
    for (int64_t t = 0; t < n_trees; ++t) {
        idx_t s = offsets[t], e = offsets[t + 1], n = e - s;
        idx_t *idx = (idx_t *)malloc((size_t)n * sizeof(idx_t));
        for (idx_t i = 0; i < n; ++i) idx[i] = i;
        partition_tree(data + s * d, idx, out + s, 0, n, 0, d, 0, target_level);
        for (idx_t i = 0; i < n; ++i) out[s + i] += s;
        free(idx);
    }
    free(data);
    return 0;
}

/* generate_rotation_matrix (pyx:576-596): python-double math, stored as fp32. */
void haet_oracle_rotation_matrix(double angle_deg, int d, float *R /* d*d row-major */)
{
    double a = angle_deg * (M_PI / 180.0); /* math.radians */
    double c = cos(a), s = sin(a);
    if (d == 2) {
        R[0] = (float)c; R[1] = (float)(-s);
        R[2] = (float)s; R[3] = (float)c;
    } else {
        R[0] = (float)(c * c);   R[1] = (float)(s * c * (s - 1)); R[2] = (float)(s * (s + c * c));
        R[3] = (float)(s * c);   R[4] = (float)(s * s * s + c * c); R[5] = (float)(s * c * (s - 1));
        R[6] = (float)(-s);      R[7] = (float)(s * c);           R[8] = (float)(c * c);
    }
}

/* leaves @ R in fp32 (pyx:646): out[i][j] = sum_k leaves[i][k]*R[k][j].  torch's CPU matmul
 * (what the reference executes here) was measured to be bit-identical to a forward FMA chain
 * acc = a0*r0; acc = fma(a_k, r_k, acc)  (4096x{2,3} random rows, 0 mismatches; the unfused and
 * the reversed orders mismatch on >50% of entries) -- see tests/test_oracle_vs_reference.py::test_rotation_matmul_is_forward_fma_chain. */
void haet_oracle_rotate(const float *leaves, int64_t n, int d, const float *R, float *out)
{
    for (int64_t i = 0; i < n; ++i)
        for (int j = 0; j < d; ++j) {
            float acc = leaves[i * d + 0] * R[0 * d + j];
            for (int k = 1; k < d; ++k) acc = fmaf(leaves[i * d + k], R[k * d + j], acc);
            out[i * d + j] = acc;
        }
}

/*
 * build_balltree_with_rotations (pyx:598-662) for trees given by offsets.
 *   rot_idx      : concatenation of the per-level index arrays
 *   rot_offsets  : n_levels+1 entries, start of each level inside rot_idx
 *   angle_deg<=0 : no rotated indices are produced (rot_offsets all 0)
 * tree_idx / tree_mask need haet_oracle_tree_slots() entries.
 */
int haet_oracle_build_with_rotations(const float *pos, int64_t n_points, int d,
                                     const int64_t *offsets, int64_t n_trees,
                                     const int32_t *strides, const int32_t *ball_sizes, int n_levels,
                                     double angle_deg,
                                     int64_t *tree_idx, uint8_t *tree_mask,
                                     int64_t *rot_idx, int64_t *rot_offsets)
{
    int rc = haet_oracle_build_balltree(pos, n_points, d, offsets, n_trees, tree_idx, tree_mask);
    if (rc) return rc;
    for (int i = 0; i <= n_levels; ++i) rot_offsets[i] = 0;
    if (!(angle_deg > 0)) return 0;

    int64_t total = haet_oracle_tree_slots(offsets, n_trees);
    /* leaves = data[tree_idx]; per-tree leaf offsets */
    float *leaves = (float *)malloc((size_t)total * d * sizeof(float));
    float *rot = (float *)malloc((size_t)total * d * sizeof(float));
    int64_t *loff = (int64_t *)malloc((size_t)(n_trees + 1) * sizeof(int64_t));
    for (int64_t i = 0; i < total; ++i)
        for (int j = 0; j < d; ++j) leaves[i * d + j] = pos[tree_idx[i] * d + j];
    loff[0] = 0;
    for (int64_t t = 0; t < n_trees; ++t)
        loff[t + 1] = loff[t] + ((idx_t)1 << tree_depth(offsets[t + 1] - offsets[t])) * 2;

    float R[9];
    haet_oracle_rotation_matrix(angle_deg, d, R);
    haet_oracle_rotate(leaves, total, d, R, rot);

    int64_t cur_n = total, w = 0;
    for (int lvl = 0; lvl < n_levels; ++lvl) {
        /* pyx:643 -- uses the TOTAL leaf count of the finest level for every level */
        double ratio = (double)total / (double)ball_sizes[lvl];
        int64_t target = (int64_t)log2(ratio); /* int() truncates toward zero */
        if (target < 0) target = 0;
        rot_offsets[lvl] = w;
        rc = haet_oracle_partition_balltree(rot, cur_n, d, loff, n_trees, target, rot_idx + w);
        if (rc) break;
        w += cur_n;
        if (lvl < n_levels - 1) {
            int s = strides[lvl];
            /* rotated.view(-1, s, d).mean(1): sequential fp32 sum then divide by s */
            int64_t nn = cur_n / s;
            for (int64_t i = 0; i < nn; ++i)
                for (int j = 0; j < d; ++j) {
                    float acc = rot[(i * s) * d + j];
                    for (int k = 1; k < s; ++k) acc = acc + rot[(i * s + k) * d + j];
                    rot[i * d + j] = acc / (float)s;
                }
            for (int64_t t = 0; t <= n_trees; ++t) loff[t] /= s;
            cur_n = nn;
        }
    }
    rot_offsets[n_levels] = w;
    free(leaves); free(rot); free(loff);
    return rc;
}
