/*
 * ggs_oracle.c -- CPU restatement of the reference rasterizer.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference fallback
 * may load this library; the product path never does (it fails loudly without the CUDA .so).
 *
 * It restates, in plain C99 with IEEE float arithmetic, the algorithm of
 * /root/reference/submodules/diff-gaussian-rasterization ("DGR/"):
 *   preprocess        DGR/cuda_rasterizer/render_forward.cu:283-386 (+ :22-78 colour, :81-243 cov2D)
 *   scan / duplicate  DGR/cuda_rasterizer/rasterizer_impl.cu:70-107, :380
 *   sort / ranges     DGR/cuda_rasterizer/rasterizer_impl.cu:37-50, :142-161, :403-421
 *   tile blend        DGR/cuda_rasterizer/render_forward.cu:391-671
 *   blend backward    DGR/cuda_rasterizer/render_backward.cu:716-1069
 *   per-Gaussian bwd  DGR/cuda_rasterizer/render_backward.cu:56-191, :193-244, :249-656, :661-713
 *   point integrate   DGR/cuda_rasterizer/sample_forward.cu:10-169, rasterizer_impl.cu:594-815
 * under the shipped compile-time configuration DGR/cuda_rasterizer/config.h:15-39.
 *
 * Parity pinning: the reference has NO golden vectors or tests for this path (SURVEY.md 4,
 * 8c).  This oracle is pinned against outputs of the reference CUDA extension itself, generated
 * on a B200 by tests/golden/make_golden.py and committed under tests/golden/ (see DESIGN.md).
 * The reference is compiled with --use_fast_math (approximate exp/div/sqrt); this file uses IEEE
 * operations, so floats agree to ~1e-6 relative and integers may differ on knife-edge inputs.
 *
 * Only the scale/rotation covariance path is restated: cov3D_precomp is unusable in the reference
 * (render_forward.cu:344-345 dereferences `scales + idx` unconditionally).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BLOCK_X 16
#define BLOCK_Y 16
#define BLOCK_SIZE 256
#define MIN_TRANSMITTANCE 0.45f
#define SPLIT 8
#define SAMPLE_RANGE 0.4f
#define SPLIT_ITERATIONS 5
#define SAMPLE_BATCH_SIZE 2

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

/* ---- glm-convention 3x3 (m.c[col][row]) --------------------------------------------------- */
typedef struct { float c[3][3]; } m3;
typedef struct { float x, y, z; } v3;

static m3 m3_cols(float a, float b, float c, float d, float e, float f, float g, float h, float i) {
    m3 m = {{{a, b, c}, {d, e, f}, {g, h, i}}};
    return m;
}
static m3 m3_mul(m3 a, m3 b) { /* glm a*b */
    m3 r;
    for (int j = 0; j < 3; j++)
        for (int i = 0; i < 3; i++)
            r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
    return r;
}
static m3 m3_t(m3 a) {
    m3 r;
    for (int j = 0; j < 3; j++)
        for (int i = 0; i < 3; i++) r.c[j][i] = a.c[i][j];
    return r;
}
static v3 m3_mulv(m3 m, v3 v) { /* glm m*v */
    v3 r = {m.c[0][0] * v.x + m.c[1][0] * v.y + m.c[2][0] * v.z, m.c[0][1] * v.x + m.c[1][1] * v.y + m.c[2][1] * v.z,
            m.c[0][2] * v.x + m.c[1][2] * v.y + m.c[2][2] * v.z};
    return r;
}
static m3 m3_outer(v3 c, v3 r) { /* glm::outerProduct(c, r): col i = c * r[i] */
    float rr[3] = {r.x, r.y, r.z};
    m3 m;
    for (int i = 0; i < 3; i++) { m.c[i][0] = c.x * rr[i]; m.c[i][1] = c.y * rr[i]; m.c[i][2] = c.z * rr[i]; }
    return m;
}
static m3 m3_scale(m3 a, float s) {
    for (int j = 0; j < 3; j++) for (int i = 0; i < 3; i++) a.c[j][i] *= s;
    return a;
}
static float v3_dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static float norm3(float x, float y, float z) { return sqrtf(x * x + y * y + z * z); }

/* auxiliary.h:51-58 */
static v3 xform43(v3 p, const float* m) {
    v3 r = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
    return r;
}

/* ---- state ------------------------------------------------------------------------------- */
typedef struct {
    int P, W, H, gx, gy, SHD, SHM, SGD, SGM, require_depth;
    int64_t D;
    float focal_x, focal_y, tanx, tany, kernel_size, scale_mod;
    float view[16], proj[16], campos[3], bg[3];
    /* GeometryState (rasterizer_impl.h:27-42) */
    float *depths, *ray_planes, *normals, *means2D, *conic_opacity, *rgb;
    uint8_t* clamped;
    int32_t* radii;
    uint32_t *tiles_touched, *point_offsets;
    /* BinningState (rasterizer_impl.h:101-110) */
    uint64_t* keys;
    uint32_t* point_list;
    /* TileState<false> / ImageState */
    uint32_t *ranges, *max_contributor, *n_contrib;
    /* point-query state */
    int64_t PN;
} Oracle;

#define FREE(p) do { free(p); (p) = NULL; } while (0)

void* ggs_oracle_create(void) { return calloc(1, sizeof(Oracle)); }

static void oracle_free_arrays(Oracle* o) {
    FREE(o->depths); FREE(o->ray_planes); FREE(o->normals); FREE(o->means2D); FREE(o->conic_opacity); FREE(o->rgb);
    FREE(o->clamped); FREE(o->radii); FREE(o->tiles_touched); FREE(o->point_offsets); FREE(o->keys);
    FREE(o->point_list); FREE(o->ranges); FREE(o->max_contributor); FREE(o->n_contrib);
}
void ggs_oracle_destroy(void* h) {
    if (!h) return;
    oracle_free_arrays((Oracle*)h);
    free(h);
}

/* rasterizer_impl.cu:37-50 */
static uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}
uint32_t ggs_oracle_higher_msb(uint32_t n) { return higher_msb(n); }

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* auxiliary.h:42-49 */
static void get_rect(float px, float py, int r, int gx, int gy, int* x0, int* y0, int* x1, int* y1) {
    *x0 = imin(gx, imax(0, (int)((px - r) / BLOCK_X)));
    *y0 = imin(gy, imax(0, (int)((py - r) / BLOCK_Y)));
    *x1 = imin(gx, imax(0, (int)((px + r + BLOCK_X - 1) / BLOCK_X)));
    *y1 = imin(gy, imax(0, (int)((py + r + BLOCK_Y - 1) / BLOCK_Y)));
}

/* Shared forward intermediates of computeCov2D (render_forward.cu:81-243) that the backward
 * (render_backward.cu:249-656) recomputes. */
typedef struct {
    v3 t;         /* clamped view-space point */
    float u, v, tc, l;
    float x_grad_mul, y_grad_mul;
    m3 J, Wm, T, R, cov, cov_cam_inv, Vrk, Vrk_inv;
    float sl[3];  /* mod * scale */
    float det0, det1, coef;
    v3 uvh, uvh_m;
} Cov2D;

static void cov2d_common(v3 mean, float fx, float fy, float tanx, float tany, float ks, const float* vm, const float* sc,
                         const float* q, float mod, Cov2D* o) {
    v3 t = xform43(mean, vm);
    o->tc = norm3(t.x, t.y, t.z);
    const float limx = 1.3f * tanx, limy = 1.3f * tany;
    float u = t.x / t.z, v = t.y / t.z;
    o->x_grad_mul = (u < -limx || u > limx) ? 0.f : 1.f;
    o->y_grad_mul = (v < -limy || v > limy) ? 0.f : 1.f;
    t.x = fminf(limx, fmaxf(-limx, u)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, v)) * t.z;
    u = t.x / t.z;
    v = t.y / t.z;
    o->t = t; o->u = u; o->v = v;
    o->J = m3_cols(fx / t.z, 0.f, -(fx * t.x) / (t.z * t.z), 0.f, fy / t.z, -(fy * t.y) / (t.z * t.z), 0.f, 0.f, 0.f);
    o->Wm = m3_cols(vm[0], vm[4], vm[8], vm[1], vm[5], vm[9], vm[2], vm[6], vm[10]);
    o->T = m3_mul(o->Wm, o->J);
    m3 S = m3_cols(1, 0, 0, 0, 1, 0, 0, 0, 1), Si = S;
    for (int k = 0; k < 3; k++) {
        o->sl[k] = mod * sc[k];
        S.c[k][k] = o->sl[k];
        Si.c[k][k] = 1.0f / o->sl[k];
    }
    float r = q[0], x = q[1], y = q[2], z = q[3];
    o->R = m3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y), 2.f * (x * y + r * z),
                   1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x), 2.f * (x * z - r * y), 2.f * (y * z + r * x),
                   1.f - 2.f * (x * x + y * y));
    m3 SR = m3_mul(S, o->R);
    m3 M = m3_mul(SR, o->T);
    o->cov = m3_mul(m3_t(M), M);
    o->Vrk = m3_mul(m3_t(SR), SR);
    m3 SiR = m3_mul(Si, o->R);
    m3 Mi = m3_mul(SiR, o->Wm);
    o->cov_cam_inv = m3_mul(m3_t(Mi), Mi);
    o->Vrk_inv = m3_mul(m3_t(SiR), SiR);
    float c00 = o->cov.c[0][0], c01 = o->cov.c[0][1], c11 = o->cov.c[1][1];
    o->det0 = fmaxf(1e-6f, c00 * c11 - c01 * c01);
    o->det1 = fmaxf(1e-6f, (c00 + ks) * (c11 + ks) - c01 * c01);
    o->coef = sqrtf(o->det0 / o->det1);
    v3 uvh = {u, v, 1.f};
    o->uvh = uvh;
    o->uvh_m = m3_mulv(o->cov_cam_inv, uvh);
    o->l = norm3(t.x, t.y, t.z);
}

/* render_forward.cu:22-78 */
static void color_from_shsg(const Oracle* o, int idx, const float* means, const float* shs, const float* sg_axis,
                            const float* sg_sharp, const float* sg_color, float* out, uint8_t* clamped) {
    float dx = means[3 * idx] - o->campos[0], dy = means[3 * idx + 1] - o->campos[1], dz = means[3 * idx + 2] - o->campos[2];
    float len = sqrtf(dx * dx + dy * dy + dz * dz);
    float x = dx / len, y = dy / len, z = dz / len;
    const float* sh = shs + (size_t)idx * o->SHM * 3;
    float res[3];
    for (int c = 0; c < 3; c++) {
        float r = SH_C0 * sh[c];
        if (o->SHD > 0) {
            r = r - SH_C1 * y * sh[3 + c] + SH_C1 * z * sh[6 + c] - SH_C1 * x * sh[9 + c];
            if (o->SHD > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                r = r + SH_C2[0] * xy * sh[12 + c] + SH_C2[1] * yz * sh[15 + c] + SH_C2[2] * (2.0f * zz - xx - yy) * sh[18 + c] +
                    SH_C2[3] * xz * sh[21 + c] + SH_C2[4] * (xx - yy) * sh[24 + c];
                if (o->SHD > 2) {
                    r = r + SH_C3[0] * y * (3.0f * xx - yy) * sh[27 + c] + SH_C3[1] * xy * z * sh[30 + c] +
                        SH_C3[2] * y * (4.0f * zz - xx - yy) * sh[33 + c] +
                        SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[36 + c] +
                        SH_C3[4] * x * (4.0f * zz - xx - yy) * sh[39 + c] + SH_C3[5] * z * (xx - yy) * sh[42 + c] +
                        SH_C3[6] * x * (xx - 3.0f * yy) * sh[45 + c];
                }
            }
        }
        res[c] = r;
    }
    for (int g = 0; g < o->SGD; g++) {
        const float* a = sg_axis + ((size_t)idx * o->SGM + g) * 3;
        float gauss = expf(sg_sharp[(size_t)idx * o->SGM + g] * ((a[0] * x + a[1] * y + a[2] * z) - 1.0f));
        const float* c = sg_color + ((size_t)idx * o->SGM + g) * 3;
        for (int k = 0; k < 3; k++) res[k] += c[k] * gauss;
    }
    for (int k = 0; k < 3; k++) {
        res[k] += 0.5f;
        clamped[3 * idx + k] = res[k] < 0;
        out[k] = fmaxf(res[k], 0.0f);
    }
}

/* render_forward.cu:283-386 */
static void preprocess_one(Oracle* o, int idx, const float* means, const float* colors_precomp, const float* opac,
                           const float* scales, const float* rots, const float* shs, const float* sg_axis,
                           const float* sg_sharp, const float* sg_color, int with_color) {
    o->radii[idx] = 0;
    o->tiles_touched[idx] = 0;
    v3 p = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
    v3 pv = xform43(p, o->view);
    if (pv.z <= 0.2f) return; /* auxiliary.h:144 */
    const float* pm = o->proj;
    float hx = pm[0] * p.x + pm[4] * p.y + pm[8] * p.z + pm[12];
    float hy = pm[1] * p.x + pm[5] * p.y + pm[9] * p.z + pm[13];
    float hw = pm[3] * p.x + pm[7] * p.y + pm[11] * p.z + pm[15];
    float pw = 1.0f / (hw + 0.0000001f);
    float projx = hx * pw, projy = hy * pw;

    Cov2D c;
    cov2d_common(p, o->focal_x, o->focal_y, o->tanx, o->tany, o->kernel_size, o->view, scales + 3 * idx, rots + 4 * idx,
                 o->scale_mod, &c);
    { /* render_forward.cu:207-241 ray plane + normal */
        float u = c.u, v = c.v, u2 = u * u, v2 = v * v, uv = u * v;
        m3 nJ_inv = m3_cols(v2 + 1, -uv, 0, -uv, u2 + 1, 0, -u, -v, 0);
        float vb = v3_dot(c.uvh_m, c.uvh);
        float ray_len2 = u2 + v2 + 1;
        float factor_normal = c.l / ray_len2;
        v3 um = {c.uvh_m.x / vb, c.uvh_m.y / vb, c.uvh_m.z / vb};
        v3 plane = m3_mulv(nJ_inv, um);
        float rsigmat = sqrtf(vb / ray_len2);
        float* rp = o->ray_planes + 4 * idx;
        rp[0] = plane.x * factor_normal / o->focal_x;
        rp[1] = plane.y * factor_normal / o->focal_y;
        rp[2] = c.tc;
        rp[3] = rsigmat;
        v3 rnv = {-plane.x * factor_normal, -plane.y * factor_normal, -1};
        m3 nJ = m3_cols(1 / c.t.z, 0.0f, -(c.t.x) / (c.t.z * c.t.z), 0.0f, 1 / c.t.z, -(c.t.y) / (c.t.z * c.t.z), c.t.x / c.l,
                        c.t.y / c.l, c.t.z / c.l);
        v3 cn = m3_mulv(nJ, rnv);
        float inv = 1.0f / sqrtf(v3_dot(cn, cn));
        o->normals[3 * idx] = cn.x * inv;
        o->normals[3 * idx + 1] = cn.y * inv;
        o->normals[3 * idx + 2] = cn.z * inv;
    }
    float cx = c.cov.c[0][0] + o->kernel_size, cy = c.cov.c[0][1], cz = c.cov.c[1][1] + o->kernel_size;
    float det = cx * cz - cy * cy;
    if (det == 0.0f) return;
    float det_inv = 1.f / det;
    float conx = cz * det_inv, cony = -cy * det_inv, conz = cx * det_inv;
    float mid = 0.5f * (cx + cz);
    float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
    float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
    float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
    /* ndc2Pix (auxiliary.h:38-40) is evaluated in double then rounded */
    float pix_x = (float)(((projx + 1.0) * o->W - 1.0) * 0.5);
    float pix_y = (float)(((projy + 1.0) * o->H - 1.0) * 0.5);
    int x0, y0, x1, y1;
    get_rect(pix_x, pix_y, (int)my_radius, o->gx, o->gy, &x0, &y0, &x1, &y1);
    if ((x1 - x0) * (y1 - y0) == 0) return;
    if (with_color) {
        if (colors_precomp == NULL)
            color_from_shsg(o, idx, means, shs, sg_axis, sg_sharp, sg_color, o->rgb + 3 * idx, o->clamped);
        else
            for (int k = 0; k < 3; k++) o->rgb[3 * idx + k] = colors_precomp[3 * idx + k];
    }
    o->depths[idx] = norm3(pv.x, pv.y, pv.z);
    o->radii[idx] = (int)my_radius;
    o->means2D[2 * idx] = pix_x;
    o->means2D[2 * idx + 1] = pix_y;
    float* co = o->conic_opacity + 4 * idx;
    co[0] = conx; co[1] = cony; co[2] = conz; co[3] = opac[idx] * c.coef;
    o->tiles_touched[idx] = (uint32_t)((y1 - y0) * (x1 - x0));
}

/* stable LSD radix sort of (key,value) on the low `bits` bits -- the semantics of
 * cub::DeviceRadixSort::SortPairs(..., 0, 32 + bit) at rasterizer_impl.cu:406-412 */
static void sort_pairs(uint64_t* keys, uint32_t* vals, int64_t n, int bits) {
    uint64_t* k2 = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(n ? n : 1));
    uint32_t* v2 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n ? n : 1));
    for (int shift = 0; shift < bits; shift += 8) {
        int64_t cnt[257] = {0};
        int w = bits - shift < 8 ? bits - shift : 8;
        uint64_t mask = (1ull << w) - 1;
        for (int64_t i = 0; i < n; i++) cnt[((keys[i] >> shift) & mask) + 1]++;
        for (int i = 0; i < 256; i++) cnt[i + 1] += cnt[i];
        for (int64_t i = 0; i < n; i++) {
            int64_t d = cnt[(keys[i] >> shift) & mask]++;
            k2[d] = keys[i];
            v2[d] = vals[i];
        }
        uint64_t* tk = keys; keys = k2; k2 = tk;
        uint32_t* tv = vals; vals = v2; v2 = tv;
    }
    /* odd number of passes leaves the result in the scratch buffers */
    if (((bits + 7) / 8) & 1) {
        memcpy(k2, keys, sizeof(uint64_t) * (size_t)n);
        memcpy(v2, vals, sizeof(uint32_t) * (size_t)n);
        uint64_t* tk = keys; keys = k2; k2 = tk;
        uint32_t* tv = vals; vals = v2; v2 = tv;
    }
    free(k2);
    free(v2);
}

/* Geometry preprocess + binning shared by forward and the point-query paths
 * (rasterizer_impl.cu:346-422, :637-714). */
static int64_t bin_gaussians(Oracle* o, const float* means, const float* colors_precomp, const float* opac,
                             const float* scales, const float* rots, const float* shs, const float* sg_axis,
                             const float* sg_sharp, const float* sg_color, int with_color) {
    const int P = o->P;
    size_t p = (size_t)(P ? P : 1);
    o->depths = (float*)calloc(p, 4);
    o->ray_planes = (float*)calloc(p * 4, 4);
    o->normals = (float*)calloc(p * 3, 4);
    o->means2D = (float*)calloc(p * 2, 4);
    o->conic_opacity = (float*)calloc(p * 4, 4);
    o->rgb = (float*)calloc(p * 3, 4);
    o->clamped = (uint8_t*)calloc(p * 3, 1);
    o->radii = (int32_t*)calloc(p, 4);
    o->tiles_touched = (uint32_t*)calloc(p, 4);
    o->point_offsets = (uint32_t*)calloc(p, 4);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++)
        preprocess_one(o, i, means, colors_precomp, opac, scales, rots, shs, sg_axis, sg_sharp, sg_color, with_color);
    uint32_t acc = 0;
    for (int i = 0; i < P; i++) { acc += o->tiles_touched[i]; o->point_offsets[i] = acc; }
    int64_t D = acc;
    o->D = D;
    o->keys = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(D ? D : 1));
    o->point_list = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(D ? D : 1));
    /* duplicateWithKeys rasterizer_impl.cu:70-107 */
#pragma omp parallel for schedule(dynamic, 1024)
    for (int i = 0; i < P; i++) {
        if (o->radii[i] > 0) {
            uint32_t off = i == 0 ? 0 : o->point_offsets[i - 1];
            int x0, y0, x1, y1;
            get_rect(o->means2D[2 * i], o->means2D[2 * i + 1], o->radii[i], o->gx, o->gy, &x0, &y0, &x1, &y1);
            uint32_t dbits;
            memcpy(&dbits, &o->depths[i], 4);
            for (int y = y0; y < y1; y++)
                for (int x = x0; x < x1; x++) {
                    uint64_t key = (uint64_t)(y * o->gx + x);
                    key <<= 32;
                    key |= dbits;
                    o->keys[off] = key;
                    o->point_list[off] = (uint32_t)i;
                    off++;
                }
        }
    }
    int bit = (int)higher_msb((uint32_t)(o->gx * o->gy));
    sort_pairs(o->keys, o->point_list, D, 32 + bit);
    int tiles = o->gx * o->gy;
    o->ranges = (uint32_t*)calloc((size_t)tiles * 2, 4);
    /* identifyTileRanges rasterizer_impl.cu:142-161 */
    for (int64_t i = 0; i < D; i++) {
        uint32_t cur = (uint32_t)(o->keys[i] >> 32);
        if (i == 0) o->ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(o->keys[i - 1] >> 32);
            if (cur != prev) { o->ranges[2 * prev + 1] = (uint32_t)i; o->ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == D - 1) o->ranges[2 * cur + 1] = (uint32_t)D;
    }
    return D;
}

static void set_camera(Oracle* o, int P, int W, int H, const float* view, const float* proj, const float* campos,
                       float tanx, float tany, float ks, float mod) {
    oracle_free_arrays(o);
    o->P = P; o->W = W; o->H = H;
    o->gx = (W + BLOCK_X - 1) / BLOCK_X;
    o->gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    o->tanx = tanx; o->tany = tany; o->kernel_size = ks; o->scale_mod = mod;
    o->focal_y = H / (2.0f * tany);
    o->focal_x = W / (2.0f * tanx);
    memcpy(o->view, view, 64);
    memcpy(o->proj, proj, 64);
    memcpy(o->campos, campos, 12);
}

/* per-pair alpha test shared by all passes (render_forward.cu:483-497) */
static inline int pair_alpha(const float* xy, const float* co, float pxf, float pyf, float* dx, float* dy, float* alpha,
                             float* G) {
    *dx = xy[0] - pxf;
    *dy = xy[1] - pyf;
    float power = -0.5f * (co[0] * *dx * *dx + co[2] * *dy * *dy) - co[1] * *dx * *dy;
    if (power > 0.0f) return 0;
    *G = expf(power);
    *alpha = fminf(0.99f, co[3] * *G);
    if (*alpha < 1.0f / 255.0f) return 0;
    return 1;
}

/* renderCUDA forward, one tile (render_forward.cu:391-671) */
static void render_tile(Oracle* o, int tile, float* out_color, float* out_alpha, float* out_normal, float* out_mdepth) {
    const int W = o->W, H = o->H;
    const int tx = tile % o->gx, ty = tile / o->gx;
    const uint32_t r0 = o->ranges[2 * tile], r1 = o->ranges[2 * tile + 1];
    const int GEOM = o->require_depth;
    float T_[BLOCK_SIZE], C_[BLOCK_SIZE][3], N_[BLOCK_SIZE][3], md0[BLOCK_SIZE];
    uint32_t last_[BLOCK_SIZE];
    uint32_t block_max = 0;
    for (int t = 0; t < BLOCK_SIZE; t++) {
        int px = tx * BLOCK_X + t % BLOCK_X, py = ty * BLOCK_Y + t / BLOCK_X;
        float T = 1.0f, C[3] = {0, 0, 0}, N[3] = {0, 0, 0}, mDepthinit = 0;
        uint32_t contributor = 0, last = 0;
        if (px < W && py < H) {
            float pxf = (float)px, pyf = (float)py;
            for (uint32_t k = r0; k < r1; k++) {
                contributor++;
                uint32_t g = o->point_list[k];
                float dx, dy, alpha, G;
                if (!pair_alpha(o->means2D + 2 * g, o->conic_opacity + 4 * g, pxf, pyf, &dx, &dy, &alpha, &G)) continue;
                float test_T = T * (1.f - alpha);
                if (test_T < 0.0001f) break;
                float aT = alpha * T;
                for (int c = 0; c < 3; c++) C[c] += o->rgb[3 * g + c] * aT;
                if (GEOM) {
                    const float* rp = o->ray_planes + 4 * g;
                    float t = rp[0] * dx + rp[1] * dy + rp[2];
                    for (int c = 0; c < 3; c++) N[c] += o->normals[3 * g + c] * aT;
                    mDepthinit = T > 0.5f ? t : mDepthinit;
                }
                T = test_T;
                last = contributor;
            }
        }
        T_[t] = T; last_[t] = last; md0[t] = mDepthinit;
        for (int c = 0; c < 3; c++) { C_[t][c] = C[c]; N_[t][c] = N[c]; }
        if (last > block_max) block_max = last;
    }
    o->max_contributor[tile] = block_max;
    for (int t = 0; t < BLOCK_SIZE; t++) {
        int px = tx * BLOCK_X + t % BLOCK_X, py = ty * BLOCK_Y + t / BLOCK_X;
        if (!(px < W && py < H)) continue;
        const int pix_id = W * py + px;
        float pxf = (float)px, pyf = (float)py;
        float T = T_[t];
        uint32_t last = last_[t];
        float mDepth = 0;
        if (GEOM) { /* render_forward.cu:549-645 */
            float T_p[SPLIT + 1];
            float depth_min = fmaxf(md0[t] - SAMPLE_RANGE, 0.f), depth_max = fmaxf(md0[t] + SAMPLE_RANGE, 0.f);
            int in_range = T <= MIN_TRANSMITTANCE;
            for (int it = 0; it < SPLIT_ITERATIONS; it++) {
                const int first = it == 0;
                const int s0 = first ? 0 : 1, s1 = first ? SPLIT + 1 : SPLIT;
                for (int s = s0; s < s1; s++) T_p[s] = 1.f;
                float interval = (depth_max - depth_min) * (1.f / SPLIT);
                int done = !in_range;
                uint32_t contributor = 0;
                for (uint32_t k = 0; k < block_max && !done; k++) {
                    contributor++;
                    done = contributor >= last;
                    uint32_t g = o->point_list[r0 + k];
                    float dx, dy, alpha, G;
                    if (!pair_alpha(o->means2D + 2 * g, o->conic_opacity + 4 * g, pxf, pyf, &dx, &dy, &alpha, &G)) continue;
                    const float* rp = o->ray_planes + 4 * g;
                    float t_peak = rp[0] * dx + rp[1] * dy + rp[2], rsigma = rp[3];
                    int ball = rsigma > 0;
                    for (int s = s0; s < s1; s++) {
                        float ts = depth_min + interval * s;
                        float delta = (ts - t_peak) * rsigma;
                        float gg = ball ? expf(-0.5f * delta * delta) : 0.f;
                        float omg = 1.f - alpha * gg;
                        float rv = 1.0f / sqrtf(omg);
                        T_p[s] *= (ts > t_peak ? (1.f - alpha) : omg) * rv;
                    }
                }
                if (first) in_range = (T_p[0] >= 0.5f) && (T_p[SPLIT] <= 0.5f) && in_range;
                int start_id = 0;
                for (int p = 1; p < SPLIT; p++) start_id = T_p[p] >= 0.5f ? p : start_id;
                depth_max = depth_min + (start_id + 1) * interval;
                depth_min = depth_min + (start_id + 0) * interval;
                float a = T_p[start_id], b = T_p[start_id + 1];
                T_p[0] = a;
                T_p[SPLIT] = b;
            }
            float w_max = (T_p[0] - 0.5f) / (T_p[0] - T_p[SPLIT]);
            w_max = !(w_max > 0.f) ? 0.f : (w_max > 1.f ? 1.f : w_max); /* __saturatef: NaN -> 0 */
            float w_min = 1.f - w_max;
            mDepth = in_range ? w_max * depth_max + w_min * depth_min : 0.f;
        }
        o->n_contrib[pix_id] = last;
        for (int c = 0; c < 3; c++) out_color[c * H * W + pix_id] = C_[t][c] + T * o->bg[c];
        out_alpha[pix_id] = 1.f - T;
        if (GEOM) {
            float pnx = (pxf - (float)(W - 1) / 2.f) / o->focal_x, pny = (pyf - (float)(H - 1) / 2.f) / o->focal_y;
            float rln = 1.0f / norm3(pnx, pny, 1.f);
            out_mdepth[pix_id] = mDepth * rln;
            float len_normal = 1.f - T;
            for (int c = 0; c < 3; c++) out_normal[c * H * W + pix_id] = last ? N_[t][c] / len_normal : 0.f;
        }
    }
}

/* Rasterizer::forward (rasterizer_impl.cu:285-448).  Outputs must be zero-initialised by the
 * caller like rasterize_points.cu:77-81 does.  Returns num_rendered. */
int64_t ggs_oracle_forward(void* h, int P, int SHD, int SHM, int SGD, int SGM, const float* bg, int W, int H,
                           const float* means, const float* colors_precomp, const float* opac, const float* scales,
                           const float* rots, const float* shs, const float* sg_axis, const float* sg_sharp,
                           const float* sg_color, float scale_mod, const float* view, const float* proj,
                           const float* campos, float tanx, float tany, float kernel_size, int require_depth,
                           float* out_color, float* out_mdepth, float* out_alpha, float* out_normal, int32_t* radii) {
    Oracle* o = (Oracle*)h;
    set_camera(o, P, W, H, view, proj, campos, tanx, tany, kernel_size, scale_mod);
    o->SHD = SHD; o->SHM = SHM; o->SGD = SGD; o->SGM = SGM; o->require_depth = require_depth;
    memcpy(o->bg, bg, 12);
    int64_t D = bin_gaussians(o, means, colors_precomp, opac, scales, rots, shs, sg_axis, sg_sharp, sg_color, 1);
    memcpy(radii, o->radii, sizeof(int32_t) * (size_t)P);
    int tiles = o->gx * o->gy;
    o->max_contributor = (uint32_t*)calloc((size_t)tiles, 4);
    o->n_contrib = (uint32_t*)calloc((size_t)W * H, 4);
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < tiles; t++) render_tile(o, t, out_color, out_alpha, out_normal, out_mdepth);
    return D;
}

/* accessors for the integer intermediates the parity tests compare bit-exactly */
#define GETTER(name, type, field) const type* ggs_oracle_##name(void* h) { return ((Oracle*)h)->field; }
GETTER(tiles_touched, uint32_t, tiles_touched)
GETTER(point_offsets, uint32_t, point_offsets)
GETTER(sorted_keys, uint64_t, keys)
GETTER(point_list, uint32_t, point_list)
GETTER(ranges, uint32_t, ranges)
GETTER(n_contrib, uint32_t, n_contrib)
GETTER(max_contributor, uint32_t, max_contributor)
GETTER(depths, float, depths)
GETTER(means2D, float, means2D)
GETTER(conic_opacity, float, conic_opacity)
GETTER(rgb, float, rgb)
GETTER(ray_planes, float, ray_planes)
GETTER(normals, float, normals)
GETTER(clamped, uint8_t, clamped)
int64_t ggs_oracle_num_rendered(void* h) { return ((Oracle*)h)->D; }

#include "ggs_oracle_backward.inc"
#include "ggs_oracle_points.inc"
#include "ggs_oracle_ncc.inc"
