// extern "C" entry points of libhaet_b200.so (see include/haet_b200.h).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace haet {

static thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};

static bool env_on(const char *name) {
    const char *e = getenv(name);
    return e != nullptr && e[0] != '\0' && !(e[0] == '0' && e[1] == '\0');
}
const Options &options() {   // function-local static: initialised once, thread-safe
    static const Options o = [] {
        Options v{};
        v.gemm_1cta = env_on("HAET_GEMM_1CTA");
        v.no_opi = env_on("HAET_NO_OPI");
        v.no_ln_fuse = env_on("HAET_NO_LN_FUSE");
        v.conv_simt = env_on("HAET_CONV_SIMT");
        v.conv_nh1 = env_on("HAET_CONV_NH1");
        v.no_pdl = env_on("HAET_NO_PDL");
        v.erwin_fused = env_on("HAET_ERWIN_FUSED");
        v.no_f32x2 = env_on("HAET_NO_F32X2");
        v.conv_xp = env_on("HAET_CONV_XP");
        v.mlp_fc2_split = env_on("HAET_MLP_FC2_SPLIT");
        auto eighths = [](const char *name, int dflt) { const char *e2 = getenv(name); const int v2 = e2 ? atoi(e2) : dflt; return v2 < 0 || v2 > 4 ? dflt : v2; };
        v.poly_p1 = eighths("HAET_POLY_P1", 1);
        v.poly_p2 = eighths("HAET_POLY_P2", 3);
        const char *e = getenv("HAET_TREE_ROLES");
        v.tree_roles = e ? atoi(e) : 0;
        e = getenv("HAET_GEMM_PF");
        v.gemm_pf = e ? atoi(e) : 2;   // measured: 0.482 -> 0.451 ms for the in-projection at distance 1-3, 0.517 at 8 (evicted before use)
        e = getenv("HAET_PEER_TIMEOUT_S");
        v.peer_timeout_s = e ? atof(e) : 60.0;
        if (!(v.peer_timeout_s > 0.0)) v.peer_timeout_s = 60.0;
        return v;
    }();
    return o;
}
Hooks &hooks() {
    static thread_local Hooks h;
    return h;
}

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static int check_shape(const haet_layer_shape *s) {
    HAET_CHECK_ARG(s != nullptr, "null shape");
    HAET_CHECK_ARG(s->B >= 0 && s->N >= 0, "negative B or N");
    HAET_CHECK_ARG(s->dim > 0 && s->heads > 0 && s->dim_head > 0 && s->slice_num > 0, "non-positive dimension");
    HAET_CHECK_ARG(s->dim_head == 4 || s->dim_head == 8 || s->dim_head == 16 || s->dim_head == 32 || s->dim_head == 64,
                   "dim_head %d not supported (4, 8, 16, 32, 64)", s->dim_head);
    HAET_CHECK_ARG(s->B * s->heads <= 65535, "B*heads > 65535 not supported");
    HAET_CHECK_ARG(s->dimensionality >= 1 && s->dimensionality <= 3 && s->dimensionality <= s->dim_head,
                   "dimensionality must be 1..3 and <= dim_head");
    HAET_CHECK_ARG(s->base_temp > 0.4f, "base_temp must exceed the 0.4 clamp (tau > 0)");
    HAET_CHECK_ARG(s->epsilon > 0.f && s->epsilon < 1.f, "epsilon must be in (0, 1)");
    return HAET_OK;
}

struct Carver {
    char *p;
    explicit Carver(void *base) : p((char *)base) {}
    template <typename T> T *take(int64_t n) {
        T *r = (T *)p;
        p += align_up(n * (int64_t)sizeof(T), 256);
        return r;
    }
};

static int64_t image_bytes(const haet_layer_shape *s) {   // packed-weight scratch of the tcgen05 GEMM
    const int inner = s->heads * s->dim_head;
    int64_t b = 4 * (int64_t)(inner > s->dim ? inner : s->dim) * (inner > s->dim ? inner : s->dim);
    if (s->conv_nd != 0 && conv_tc_supported(s) && conv_tc_image_bytes(s) > b) b = conv_tc_image_bytes(s);
    return align_up(b, 256);
}
static int64_t deslice_o_bytes(const haet_layer_shape *s) {   // fp32 o, or its operand image (rows padded to 256)
    return align_up(opi_bytes(s->B * s->N, s->heads * s->dim_head), 256);
}
// pass 2 writes o as the to_out GEMM's operand image when both run on the tensor-core path
static bool out_opi(const haet_layer_shape *s) {
    return hooks().path == 0 && slice_tc_supported(s) && deslice_tc_supported(s) && gemm_opi_supported(s->heads * s->dim_head, s->dim);
}
static int64_t deslice_ws_bytes(const haet_layer_shape *s) { return deslice_o_bytes(s) + image_bytes(s); }
static int64_t slice_parts_bytes(const haet_layer_shape *s) {
    int64_t c = slice_chunks(s);
    if (slice_tc_supported(s) && slice_tc_chunks(s) > c) c = slice_tc_chunks(s);
    return align_up(c * haet_partial_floats(s) * (int64_t)sizeof(float), 256);
}

struct LayerBuf {
    void *spill; void *ws; float *partial, *eidetic, *lse2, *pos, *states; int64_t *tree_idx; uint8_t *tree_mask;
    int64_t *rot[HAET_MAX_LEVELS]; void *erwin_ws; float *xdev, *odev; int64_t slots; int64_t total;
};

static int carve_layer(const haet_layer_shape *s, const haet_erwin_cfg *c, void *base, bool host_io, LayerBuf &b) {
    const int64_t T = s->B * s->heads * s->slice_num;
    b.slots = haet_balltree_slots(s->B * s->heads, s->slice_num);
    HAET_CHECK_ARG(b.slots >= 0, "slice_num must be >= 2");
    Carver cv(base);
    b.spill = cv.take<char>(haet_spill_bytes(s));
    int64_t w1 = haet_slice_workspace_bytes(s), w2 = deslice_ws_bytes(s);
    b.ws = cv.take<char>(w1 > w2 ? w1 : w2);
    b.partial = cv.take<float>(haet_partial_floats(s));
    b.eidetic = cv.take<float>(T * s->dim_head);
    b.lse2 = cv.take<float>(T);
    b.pos = cv.take<float>(T * 3);
    b.states = cv.take<float>(T * s->dim_head);
    b.tree_idx = cv.take<int64_t>(b.slots);
    b.tree_mask = cv.take<uint8_t>(b.slots);
    for (int i = 0; i < HAET_MAX_LEVELS; ++i) b.rot[i] = i < c->n_levels ? cv.take<int64_t>(b.slots) : nullptr;
    int64_t ew = haet_erwin_workspace_bytes(c, b.slots);
    HAET_CHECK_ARG(ew >= 0, "bad erwin config");
    b.erwin_ws = cv.take<char>(ew);
    b.xdev = b.odev = nullptr;
    if (host_io) {
        b.xdev = cv.take<float>(s->B * s->N * (int64_t)s->dim);
        b.odev = cv.take<float>(s->B * s->N * (int64_t)s->dim);
    }
    b.total = cv.p - (char *)base;
    return HAET_OK;
}

}  // namespace haet

using namespace haet;

extern "C" {

int haet_version(void) { return 3; }   // 3: haet_erwin_cfg gained bn_train / bn_eps / bn_stats_d
const char *haet_last_error(void) { return g_err; }
int64_t haet_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int haet_device_check(void) {
    int dev = 0;
    cudaDeviceProp p;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&p, dev) != cudaSuccess) {
        set_error("no CUDA device");
        return HAET_ENODEV;
    }
    if (p.major != 10) {
        set_error("device %s is sm_%d%d; libhaet_b200 is built for sm_100a only", p.name, p.major, p.minor);
        return HAET_ENODEV;
    }
    return HAET_OK;
}

int64_t haet_spill_bytes(const haet_layer_shape *s) {
    SpillView v = spill_view(s, nullptr);
    int64_t bytes = (char *)v.xp - (char *)nullptr;
    if (spill_needs_xp(s)) bytes += align_up(s->B * s->N * (int64_t)s->heads * s->dim_head * (int64_t)sizeof(float), 256);
    return bytes;
}
int64_t haet_partial_floats(const haet_layer_shape *s) {
    return s->B * s->heads * (int64_t)s->slice_num * (s->dim_head + 2);
}
int64_t haet_slice_workspace_bytes(const haet_layer_shape *s) { return slice_parts_bytes(s) + image_bytes(s); }
int haet_set_path(int path) {
    if (path != 0 && path != 1) { set_error("path must be 0 (tensor core) or 1 (fp32 SIMT)"); return HAET_EINVAL; }
    hooks().path = path;
    return HAET_OK;
}
int haet_set_stage_mask(int mask) {
    if (mask < 0 || mask > 15) { set_error("stage mask must be in 0..15"); return HAET_EINVAL; }
    hooks().stage_mask = mask;
    return HAET_OK;
}
int haet_set_raise_threshold(float log2_units) {
    if (!(log2_units >= 0.f) || log2_units > 15.f) { set_error("raise threshold must be in [0, 15] log2 units (0 = default)"); return HAET_EINVAL; }
    hooks().st_raise = log2_units;
    return HAET_OK;
}
int haet_debug_gemm(const float *x_d, const float *w_d, const float *bias_d, float *y_d, int64_t M, int32_t K, int32_t Nn,
                    void *img_d, void *stream) {
    if (hooks().path == 0 && gemm_tc_supported(K, Nn)) return launch_gemm_tc(x_d, w_d, bias_d, y_d, M, K, Nn, img_d, (cudaStream_t)stream);
    return launch_linear(x_d, w_d, bias_d, y_d, M, K, Nn, (cudaStream_t)stream);
}
int64_t haet_deslice_workspace_bytes(const haet_layer_shape *s) { return deslice_ws_bytes(s); }

// ln_1 scratch behind the regular slice workspace: folded weights + row statistics when the in-projection fuses it, else normalised x
static bool ln_fuse_ok(const haet_layer_shape *s) {
    return hooks().path == 0 && slice_tc_supported(s) && deslice_tc_supported(s) && s->conv_nd == 0 && s->dim_head != 64 &&
           gemm_tc_supported(s->dim, s->heads * s->dim_head) && ln_fused_supported(s);
}
int64_t haet_slice_ln_workspace_bytes(const haet_layer_shape *s) {
    const int64_t extra = ln_fuse_ok(s) ? ln_fused_scratch_bytes(s) : align_up(s->B * s->N * (int64_t)s->dim * 4, 256);
    return align_up(haet_slice_workspace_bytes(s), 256) + extra;
}

static int slice_reduce_impl(const haet_layer_shape *s, const float *x_d, const float *wx_d, const float *bx_d,
                             const float *ws_d, const float *bs_d, const float *wt_d, float bt, void *spill_d,
                             float *partial_d, void *workspace_d, void *stream, const float *ln_g_d, const float *ln_b_d,
                             float ln_eps);

int haet_slice_reduce(const haet_layer_shape *s, const float *x_d, const float *wx_d, const float *bx_d,
                      const float *ws_d, const float *bs_d, const float *wt_d, float bt, void *spill_d,
                      float *partial_d, void *workspace_d, void *stream) {
    return slice_reduce_impl(s, x_d, wx_d, bx_d, ws_d, bs_d, wt_d, bt, spill_d, partial_d, workspace_d, stream, nullptr, nullptr, 0.f);
}

int haet_slice_reduce_ln(const haet_layer_shape *s, const float *x_d, const float *ln_g_d, const float *ln_b_d, float ln_eps,
                         const float *wx_d, const float *bx_d, const float *ws_d, const float *bs_d, const float *wt_d,
                         float bt, void *spill_d, float *partial_d, void *workspace_d, void *stream) {
    HAET_CHECK_ARG(ln_g_d != nullptr && ln_b_d != nullptr, "haet_slice_reduce_ln: LayerNorm weight and bias are required");
    HAET_CHECK_ARG(ln_eps > 0.f, "haet_slice_reduce_ln: eps must be positive");
    return slice_reduce_impl(s, x_d, wx_d, bx_d, ws_d, bs_d, wt_d, bt, spill_d, partial_d, workspace_d, stream, ln_g_d, ln_b_d, ln_eps);
}

static int slice_reduce_impl(const haet_layer_shape *s, const float *x_d, const float *wx_d, const float *bx_d,
                             const float *ws_d, const float *bs_d, const float *wt_d, float bt, void *spill_d,
                             float *partial_d, void *workspace_d, void *stream, const float *ln_g_d, const float *ln_b_d,
                             float ln_eps) {
    int rc = check_shape(s);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const SpillView sv = spill_view(s, spill_d);
    float *xp = sv.xp;
    if (s->B == 0) return HAET_OK;
    HAET_CHECK_ARG(s->N > 0, "a rank needs at least one point (N == 0)");
    float *parts = (float *)workspace_d;
    void *img = (char *)workspace_d + slice_parts_bytes(s);
    const int inner = s->heads * s->dim_head;
    const bool tc = hooks().path == 0 && slice_tc_supported(s) && deslice_tc_supported(s);
    LnArgs ln{ln_g_d, ln_b_d, ln_eps, (char *)workspace_d + align_up(haet_slice_workspace_bytes(s), 256)};
    const bool ln_fused = ln_g_d != nullptr && ln_fuse_ok(s);
    if (ln_g_d != nullptr && !ln_fused && (hooks().stage_mask & 1)) {
        // shapes / paths without the fused kernel: normalise into the scratch, then the regular route
        float *xn = (float *)ln.scratch;
        if ((rc = launch_layernorm_rows(x_d, ln_g_d, ln_b_d, ln_eps, xn, s->B * s->N, s->dim, st))) return rc;
        x_d = xn;
    }
    if (!(hooks().stage_mask & 1)) {
        rc = HAET_OK;   // bench hook: the spill of an earlier call is reused
    } else if (tc && s->conv_nd == 0 && gemm_tc_supported(s->dim, inner) && s->dim_head != 64) {
        // Linear front-end: the GEMM epilogue writes the operand image directly
        rc = launch_inproj_tc(s, x_d, wx_d, bx_d, wt_d, bt, sv.img, sv.scale, img, st, ln_fused ? &ln : nullptr);
    } else {
        bool packed = false;
        if (hooks().path == 0 && s->conv_nd == 0 && gemm_tc_supported(s->dim, inner))
            rc = launch_gemm_tc(x_d, wx_d, bx_d, xp, s->B * s->N, s->dim, inner, img, st);
        else if (tc && s->conv_nd != 0 && conv_tc_supported(s) && !options().conv_simt && !options().conv_xp && (s->dim_head == 32 || s->dim_head == 16)) {
            // Conv2d / Conv3d front-end: implicit GEMM on tcgen05 whose epilogue writes the operand image itself
            rc = launch_conv_tc_img(s, x_d, wx_d, bx_d, wt_d, bt, sv.img, sv.scale, img, st);
            packed = true;
        } else if (hooks().path == 0 && s->conv_nd != 0 && conv_tc_supported(s) && !options().conv_simt)
            rc = launch_conv_tc(s, x_d, wx_d, bx_d, xp, img, st);   // the same to fp32 x_proj (SIMT passes, dim_head 64: packed below)
        else
            rc = launch_in_project(s, x_d, wx_d, bx_d, xp, st);
        if (!rc && tc && !packed) rc = launch_xpi_pack(s, xp, wt_d, bt, sv.img, sv.scale, st);
    }
    if (rc) return rc;
    if (!(hooks().stage_mask & 2)) return HAET_OK;
    if (tc) {
        if ((rc = launch_slice_tc(s, sv.img, sv.scale, ws_d, bs_d, parts, st))) return rc;
        return launch_slice_merge(s, parts, slice_tc_chunks(s), partial_d, nullptr, nullptr, st);
    }
    if ((rc = launch_slice_partial(s, xp, ws_d, bs_d, wt_d, bt, parts, st))) return rc;
    return launch_slice_merge(s, parts, slice_chunks(s), partial_d, nullptr, nullptr, st);
}

// ---- halo-exchange sharding of the conv front-ends: a rank owns a slab of grid planes ----------------------
static int halo_ext_shape(const haet_layer_shape *s, int32_t halo_lo, int32_t halo_hi, haet_layer_shape *ext) {
    int rc = check_shape(s);
    if (rc) return rc;
    HAET_CHECK_ARG(s->conv_nd == 2 || s->conv_nd == 3, "haet_slice_reduce_halo: the shape has no conv front-end (conv_nd must be 2 or 3)");
    const int64_t plane = (int64_t)s->grid[1] * (s->conv_nd == 3 ? s->grid[2] : 1);
    HAET_CHECK_ARG(plane > 0 && s->grid[0] > 0 && s->N == (int64_t)s->grid[0] * plane,
                   "haet_slice_reduce_halo: N = %lld is not grid[0] * grid[1] (* grid[2]) of the local slab", (long long)s->N);
    const int r = s->kernel / 2;
    HAET_CHECK_ARG((halo_lo == 0 || halo_lo == r) && (halo_hi == 0 || halo_hi == r),
                   "haet_slice_reduce_halo: a halo is 0 (outer boundary of the grid) or kernel / 2 = %d planes", r);
    *ext = *s;
    ext->grid[0] = s->grid[0] + halo_lo + halo_hi;
    ext->N = (int64_t)ext->grid[0] * plane;
    return HAET_OK;
}
static bool halo_tc_ok(const haet_layer_shape *s) { return hooks().path == 0 && slice_tc_supported(s) && deslice_tc_supported(s); }

int64_t haet_slice_halo_workspace_bytes(const haet_layer_shape *s, int32_t halo_lo, int32_t halo_hi, int32_t with_ln) {
    haet_layer_shape ext;
    if (halo_ext_shape(s, halo_lo, halo_hi, &ext) != HAET_OK) return HAET_EINVAL;
    if (!halo_tc_ok(s)) { set_error("haet_slice_reduce_halo needs the tensor-core passes (dim_head 16 / 32 / 64)"); return HAET_ENOSUP; }
    const int64_t ib = image_bytes(s) > image_bytes(&ext) ? image_bytes(s) : image_bytes(&ext);
    const int64_t inner = (int64_t)s->heads * s->dim_head;
    return align_up(slice_parts_bytes(s) + ib, 256) + align_up(s->B * ext.N * inner * 4, 256) +
           (with_ln ? align_up(s->B * ext.N * (int64_t)s->dim * 4, 256) : 0);
}

int haet_slice_reduce_halo(const haet_layer_shape *s, int32_t halo_lo, int32_t halo_hi, const float *x_ext_d, const float *ln_g_d,
                           const float *ln_b_d, float ln_eps, const float *wx_d, const float *bx_d, const float *ws_d,
                           const float *bs_d, const float *wt_d, float bt, void *spill_d, float *partial_d, void *workspace_d,
                           void *stream) {
    haet_layer_shape ext;
    int rc = halo_ext_shape(s, halo_lo, halo_hi, &ext);
    if (rc) return rc;
    if (!halo_tc_ok(s)) { set_error("haet_slice_reduce_halo needs the tensor-core passes (dim_head 16 / 32 / 64)"); return HAET_ENOSUP; }
    if (s->B == 0) return HAET_OK;
    HAET_CHECK_ARG((ln_g_d == nullptr) == (ln_b_d == nullptr) && (ln_g_d == nullptr || ln_eps > 0.f), "haet_slice_reduce_halo: bad LayerNorm arguments");
    cudaStream_t st = (cudaStream_t)stream;
    const SpillView sv = spill_view(s, spill_d);
    const int64_t inner = (int64_t)s->heads * s->dim_head, plane = s->N / s->grid[0];
    const int64_t ib = image_bytes(s) > image_bytes(&ext) ? image_bytes(s) : image_bytes(&ext);
    float *parts = (float *)workspace_d;
    void *img = (char *)workspace_d + slice_parts_bytes(s);
    float *xp_ext = (float *)((char *)workspace_d + align_up(slice_parts_bytes(s) + ib, 256));
    if (hooks().stage_mask & 1) {
        const float *xin = x_ext_d;
        if (ln_g_d) {   // the caller's ln_1 on every row of the extended slab (zero padding applies AFTER the LayerNorm, as in the reference)
            float *xn = (float *)((char *)xp_ext + align_up(s->B * ext.N * inner * 4, 256));
            if ((rc = launch_layernorm_rows(x_ext_d, ln_g_d, ln_b_d, ln_eps, xn, s->B * ext.N, s->dim, st))) return rc;
            xin = xn;
        }
        // the conv over the slab PLUS its halo planes: the zero padding beyond the extended slab is the true padding at an outer
        // boundary (halo 0) and only reaches halo rows otherwise, which the image pack below skips
        if (conv_tc_supported(&ext) && !options().conv_simt) rc = launch_conv_tc(&ext, xin, wx_d, bx_d, xp_ext, img, st);
        else rc = launch_in_project(&ext, xin, wx_d, bx_d, xp_ext, st);
        if (rc) return rc;
        if ((rc = launch_xpi_pack(s, xp_ext + (int64_t)halo_lo * plane * inner, wt_d, bt, sv.img, sv.scale, st, ext.N))) return rc;
    }
    if (!(hooks().stage_mask & 2)) return HAET_OK;
    if ((rc = launch_slice_tc(s, sv.img, sv.scale, ws_d, bs_d, parts, st))) return rc;
    return launch_slice_merge(s, parts, slice_tc_chunks(s), partial_d, nullptr, nullptr, st);
}

int haet_slice_merge(const haet_layer_shape *s, const float *partials_d, int32_t n_ranks, float *eidetic_d,
                     float *lse2_d, float *pos_d, void *stream) {
    int rc = check_shape(s);
    if (rc) return rc;
    HAET_CHECK_ARG(n_ranks >= 1, "n_ranks must be >= 1");
    if (s->B == 0) return HAET_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if ((rc = launch_slice_merge(s, partials_d, n_ranks, nullptr, eidetic_d, lse2_d, st))) return rc;
    return launch_slice_pos(s, eidetic_d, pos_d, st);
}

int haet_deslice_out(const haet_layer_shape *s, const void *spill_d, const float *lse2_d, const float *states_d,
                     const float *ws_d, const float *bs_d, const float *wt_d, float bt, const float *wo_d,
                     const float *bo_d, float *out_d, void *workspace_d, void *stream) {
    return haet_deslice_out_res(s, spill_d, lse2_d, states_d, ws_d, bs_d, wt_d, bt, wo_d, bo_d, nullptr, out_d, workspace_d, stream);
}

int haet_deslice_out_res(const haet_layer_shape *s, const void *spill_d, const float *lse2_d, const float *states_d,
                         const float *ws_d, const float *bs_d, const float *wt_d, float bt, const float *wo_d,
                         const float *bo_d, const float *resid_d, float *out_d, void *workspace_d, void *stream) {
    int rc = check_shape(s);
    if (rc) return rc;
    if (s->B == 0 || s->N == 0) return HAET_OK;
    cudaStream_t st = (cudaStream_t)stream;
    float *o = (float *)workspace_d;
    const SpillView sv = spill_view(s, const_cast<void *>(spill_d));
    if (!(hooks().stage_mask & 4))
        rc = HAET_OK;
    else if (hooks().path == 0 && slice_tc_supported(s) && deslice_tc_supported(s))
        rc = launch_deslice_tc(s, sv.img, sv.scale, lse2_d, states_d, ws_d, bs_d, o, out_opi(s) ? 1 : 0, st);
    else
        rc = launch_deslice(s, sv.xp, lse2_d, states_d, ws_d, bs_d, wt_d, bt, o, st);
    if (rc) return rc;
    if (!(hooks().stage_mask & 8)) return HAET_OK;
    const int inner = s->heads * s->dim_head;
    HAET_CHECK_ARG(resid_d != out_d || resid_d == nullptr, "the residual must not alias the output");
    if (out_opi(s))
        return launch_gemm_tc_opi((const uint8_t *)o, wo_d, bo_d, out_d, s->B * s->N, inner, s->dim,
                                  (char *)workspace_d + deslice_o_bytes(s), st, resid_d);
    if (hooks().path == 0 && gemm_tc_supported(inner, s->dim))
        return launch_gemm_tc(o, wo_d, bo_d, out_d, s->B * s->N, inner, s->dim, (char *)workspace_d + deslice_o_bytes(s), st, resid_d);
    if ((rc = launch_linear(o, wo_d, bo_d, out_d, s->B * s->N, inner, s->dim, st))) return rc;
    return resid_d ? launch_add_rows(out_d, resid_d, s->B * s->N * (int64_t)s->dim, st) : HAET_OK;
}

int64_t haet_block_mlp_workspace_bytes(int64_t M, int32_t dim, int32_t hidden) {
    if (M < 0 || !mlp_tc_supported(dim, hidden)) {
        set_error("haet_block_mlp: unsupported dim=%d hidden=%d (dim 256, hidden a multiple of 256, tensor-core path)", dim, hidden);
        return HAET_ENOSUP;
    }
    return mlp_tc_workspace_bytes(M, dim, hidden);
}

int haet_block_mlp(const float *x_d, int64_t M, int32_t dim, int32_t hidden, const float *ln_g_d, const float *ln_b_d,
                   float ln_eps, const float *w1_d, const float *b1_d, const float *w2_d, const float *b2_d, float *out_d,
                   void *workspace_d, void *stream) {
    HAET_CHECK_ARG(M >= 0 && x_d && ln_g_d && ln_b_d && w1_d && b1_d && w2_d && b2_d && out_d && workspace_d, "haet_block_mlp: null argument");
    HAET_CHECK_ARG(ln_eps > 0.f, "haet_block_mlp: eps must be positive");
    HAET_CHECK_ARG(x_d != out_d, "haet_block_mlp: the output must not alias the input");
    if (!mlp_tc_supported(dim, hidden)) {
        set_error("haet_block_mlp: unsupported dim=%d hidden=%d (dim 256, hidden a multiple of 256, tensor-core path)", dim, hidden);
        return HAET_ENOSUP;
    }
    return launch_mlp_tc(x_d, M, dim, hidden, ln_g_d, ln_b_d, ln_eps, w1_d, b1_d, w2_d, b2_d, out_d, workspace_d, (cudaStream_t)stream);
}

int64_t haet_layer_scratch_bytes(const haet_layer_shape *s, const haet_erwin_cfg *c) {
    LayerBuf b;
    if (check_shape(s) || carve_layer(s, c, nullptr, true, b)) return -1;
    return b.total;
}

static int layer_forward_impl(const haet_layer_shape *s, const haet_erwin_cfg *c, const haet_layer_weights *w,
                              const float *x_d, float *out_d, LayerBuf &b, cudaStream_t st) {
    int rc;
    if ((rc = haet_slice_reduce(s, x_d, w->wx_d, w->bx_d, w->ws_d, w->bs_d, w->wt_d, w->bt, b.spill, b.partial, b.ws, st))) return rc;
    if ((rc = haet_slice_merge(s, b.partial, 1, b.eidetic, b.lse2, b.pos, st))) return rc;
    if (s->B == 0) return HAET_OK;
    const int64_t T = s->B * s->heads * s->slice_num;
    if ((rc = haet_balltree_build(b.pos, s->B * s->heads, s->slice_num, s->dimensionality, c->strides, c->ball_sizes,
                                  c->n_levels, (double)c->rotate, b.tree_idx, b.tree_mask, b.rot, st))) return rc;
    if ((rc = haet_erwin_forward_trees(c, b.eidetic, b.pos, T, s->B * s->heads, b.tree_idx, b.tree_mask, b.slots, b.rot, w->erwin_d,
                                       b.erwin_ws, b.states, st))) return rc;
    return haet_deslice_out(s, b.spill, b.lse2, b.states, w->ws_d, w->bs_d, w->wt_d, w->bt, w->wo_d, w->bo_d, out_d, b.ws, st);
}

int haet_layer_forward(const haet_layer_shape *s, const haet_erwin_cfg *c, const haet_layer_weights *w,
                       const float *x_d, float *out_d, void *scratch_d, void *stream) {
    int rc = check_shape(s);
    if (rc) return rc;
    HAET_CHECK_ARG(c->dimensionality == s->dimensionality && c->c_hidden[0] == s->dim_head,
                   "erwin config does not match the layer (dimensionality / c_hidden[0])");
    LayerBuf b;
    if ((rc = carve_layer(s, c, scratch_d, false, b))) return rc;
    return layer_forward_impl(s, c, w, x_d, out_d, b, (cudaStream_t)stream);
}

int haet_layer_forward_host(const haet_layer_shape *s, const haet_erwin_cfg *c, const haet_layer_weights *w,
                            const float *x_h, float *out_h, void *scratch_d, void *stream) {
    int rc = check_shape(s);
    if (rc) return rc;
    HAET_CHECK_ARG(c->dimensionality == s->dimensionality && c->c_hidden[0] == s->dim_head,
                   "erwin config does not match the layer (dimensionality / c_hidden[0])");
    LayerBuf b;
    if ((rc = carve_layer(s, c, scratch_d, true, b))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t bytes = sizeof(float) * (size_t)(s->B * s->N) * s->dim;
    if (bytes == 0) return HAET_OK;
    HAET_CUDA(cudaMemcpyAsync(b.xdev, x_h, bytes, cudaMemcpyHostToDevice, st));
    if ((rc = layer_forward_impl(s, c, w, b.xdev, b.odev, b, st))) return rc;
    HAET_CUDA(cudaMemcpyAsync(out_h, b.odev, bytes, cudaMemcpyDeviceToHost, st));
    return HAET_OK;
}

}  // extern "C"
