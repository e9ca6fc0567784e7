/*
 * kbmod_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C, gcc) of the reference algorithm
 * for KBMOD's shift-and-stack hot path.  Nothing in the product path (kbmod_b200/) may call this; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg do, and only as the checker.
 *
 * Parity is PINNED: tests/test_oracle_pins.py checks every function here against (a) the known-answer
 * vectors in the reference's own tests and (b) golden fixtures produced by the reference itself
 * (oracle/_ref built from /root/reference by oracle/build_ref.sh; generator tests/golden/make_golden.py).
 *
 * Build: gcc -O2 -std=c11 -fopenmp -ffp-contract=off -shared -fPIC kbmod_oracle.c -lm
 * (-ffp-contract=off: the reference's host build has no FMA contraction; every float op rounds once.)
 *
 * Citations are file:line under /root/reference/src/kbmod/search/.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

static inline int valid(float v) { return isfinite(v); } /* common.h:41 pixel_value_valid */

/* ---- P2: mask-aware, renormalised PSF correlation (image_utils_cpp.cpp:20-68) ------------------- */
ORC_API void orc_convolve(const float *img, int H, int W, const float *psf, int P, float *out) {
    const int rad = (P - 1) / 2;
    float psf_total = 0.0f;
    for (int r = 0; r < P; ++r)
        for (int c = 0; c < P; ++c) psf_total += psf[r * P + c]; /* :30-35 row-major f32 adds */
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y) {
        for (int x = 0; x < W; ++x) {
            const float centre = img[(size_t)y * W + x];
            if (!valid(centre)) { /* :41-44 masked pixels pass through */
                out[(size_t)y * W + x] = centre;
                continue;
            }
            float sum = 0.0f, portion = 0.0f;
            for (int j = -rad; j <= rad; ++j) {     /* :48 j outer */
                for (int i = -rad; i <= rad; ++i) { /* :49 i inner */
                    const int xx = x + i, yy = y + j;
                    if (xx < 0 || xx >= W || yy < 0 || yy >= H) continue;
                    const float pix = img[(size_t)yy * W + xx];
                    if (!valid(pix)) continue;
                    const float w = psf[(j + rad) * P + (i + rad)];
                    portion += w;
                    sum += pix * w; /* separate mul and add */
                }
            }
            out[(size_t)y * W + x] = (portion == 0.0f) ? NAN : (sum * psf_total) / portion; /* :60-64 */
        }
    }
}

/* ---- P1 / P1': psi = conv(sci/var, psf), phi = conv(1/var, psf^2) (image_utils_cpp.cpp:110-177) --- */
ORC_API void orc_generate_psi(const float *sci, const float *var, int H, int W, const float *psf, int P,
                              float *out) {
    const size_t n = (size_t)H * W;
    float *tmp = (float *)malloc(n * sizeof(float));
    for (size_t p = 0; p < n; ++p) {
        const float v = var[p];
        tmp[p] = (isfinite(v) && v != 0.0f && isfinite(sci[p])) ? sci[p] / v : NAN; /* :142-149 */
    }
    orc_convolve(tmp, H, W, psf, P, out);
    free(tmp);
}

ORC_API void orc_generate_phi(const float *var, int H, int W, const float *psf, int P, float *out) {
    const size_t n = (size_t)H * W;
    float *tmp = (float *)malloc(n * sizeof(float));
    float *psf2 = (float *)malloc((size_t)P * P * sizeof(float));
    for (size_t p = 0; p < n; ++p) {
        const float v = var[p];
        tmp[p] = (isfinite(v) && v != 0.0f) ? (float)(1.0 / (double)v) : NAN; /* :166-171 double div */
    }
    for (int k = 0; k < P * P; ++k) psf2[k] = psf[k] * psf[k]; /* :110-120 */
    orc_convolve(tmp, H, W, psf2, P, out);
    free(tmp);
    free(psf2);
}

/* ---- P5: scale parameters (psi_phi_array.cpp:219-245) -------------------------------------------- */
ORC_API void orc_scale_params(const float *vals, uint64_t n, int num_bytes, float out3[3]) {
    float mn = FLT_MAX, mx = -FLT_MAX;
    for (uint64_t i = 0; i < n; ++i) {
        const float e = vals[i];
        if (valid(e)) {
            mn = (e < mn) ? e : mn; /* std::min(min_val, elem) */
            mx = (mx < e) ? e : mx; /* std::max(max_val, elem) */
        }
    }
    float scale = 1.0f;
    if (num_bytes == 1 || num_bytes == 2) {
        float width = mx - mn;
        if (width < 1e-6) width = 1e-6; /* :237 compares and assigns through double 1e-6 */
        const uint64_t num_values = (1u << (8 * num_bytes)) - 1;
        scale = (float)((double)width / (double)num_values);
    }
    out3[0] = mn; out3[1] = mx; out3[2] = scale;
}

/* ---- P6: encode (psi_phi_array_ds.h:40-43, psi_phi_array.cpp:247-291) ----------------------------- */
static inline float encode_scalar(float v, float mn, float safe_mx, float scale) {
    if (!valid(v)) return 0.0f;
    float c = (v < safe_mx) ? v : safe_mx; /* std::min(value, max_val) */
    c = (c < mn) ? mn : c;                 /* std::max(., min_val)     */
    return (float)((double)((c - mn) / scale) + 1.0);
}

/* psi, phi: [n_pix_total] planes; out: interleaved codes, 1 or 2 bytes each; params = {min,max,scale}. */
ORC_API void orc_encode(const float *psi, const float *phi, uint64_t n, int num_bytes, const float psi_p[3],
                        const float phi_p[3], void *out) {
    const float safe_psi = (float)((double)psi_p[1] - (double)psi_p[2] / 100.0); /* :264 */
    const float safe_phi = (float)((double)phi_p[1] - (double)phi_p[2] / 100.0); /* :265 */
    for (uint64_t i = 0; i < n; ++i) {
        const float a = encode_scalar(psi[i], psi_p[0], safe_psi, psi_p[2]);
        const float b = encode_scalar(phi[i], phi_p[0], safe_phi, phi_p[2]);
        if (num_bytes == 1) {
            ((uint8_t *)out)[2 * i] = (uint8_t)(int)a; /* :284-285 static_cast<T>: truncate, wrap */
            ((uint8_t *)out)[2 * i + 1] = (uint8_t)(int)b;
        } else {
            ((uint16_t *)out)[2 * i] = (uint16_t)(int)a;
            ((uint16_t *)out)[2 * i + 1] = (uint16_t)(int)b;
        }
    }
}

/* ---- stack description shared by the search functions ------------------------------------------- */
typedef struct {
    const void *data;  /* reference layout [t][row][col][psi,phi] */
    const double *times;
    int T, H, W, num_bytes; /* num_bytes 4 (float), 2 (uint16), 1 (uint8) */
    float psi_min, psi_scale, phi_min, phi_scale;
} orc_stack;

/* P8: read_encoded_psi_phi (kernels/kernels.cu:37-71) */
static inline void read_pp(const orc_stack *s, int t, int row, int col, float *psi, float *phi) {
    if (row < 0 || col < 0 || row >= s->H || col >= s->W) { *psi = NAN; *phi = NAN; return; }
    const uint64_t idx = 2 * ((uint64_t)s->H * s->W * (uint64_t)t + (uint64_t)row * s->W + (uint64_t)col);
    if (s->num_bytes == 4) {
        *psi = ((const float *)s->data)[idx];
        *phi = ((const float *)s->data)[idx + 1];
        return;
    }
    const float a = (s->num_bytes == 1) ? (float)((const uint8_t *)s->data)[idx] : (float)((const uint16_t *)s->data)[idx];
    const float b = (s->num_bytes == 1) ? (float)((const uint8_t *)s->data)[idx + 1] : (float)((const uint16_t *)s->data)[idx + 1];
    *psi = (a == 0.0f) ? NAN : (float)(((double)a - 1.0) * (double)s->psi_scale + (double)s->psi_min);
    *phi = (b == 0.0f) ? NAN : (float)(((double)b - 1.0) * (double)s->phi_scale + (double)s->phi_min);
}

ORC_API void orc_read_psi_phi(const orc_stack *s, int t, int row, int col, float out[2]) {
    read_pp(s, t, row, col, &out[0], &out[1]);
}

/* S5: predict_index (kernels.cu:33-35): fp64, no contraction on the host build. */
static inline int predict_index(float pos0, float vel0, double time) {
    const double prod = (double)vel0 * time;
    const double s = (double)pos0 + prod;
    return (int)floor(s + 0.5);
}

/* ---- S4: sigma-G keep range over sorted indices (kernels.cu:77-147) -------------------------------- */
ORC_API void orc_sigmag_indices(const float *values, int n, float sgl0, float sgl1, float coeff, float width,
                                int *idx, int *min_keep, int *max_keep) {
    if (n == 0) { *min_keep = 0; *max_keep = -1; return; }
    if (sgl0 < 0.0001) sgl0 = 0.0001;
    if (sgl1 > 0.9999) sgl1 = 0.9999;
    for (int j = 0; j < n; ++j) idx[j] = j;
    for (int j = 0; j < n; ++j)          /* exchange sort, same visiting order => same tie order */
        for (int k = j + 1; k < n; ++k)
            if (values[idx[j]] > values[idx[k]]) { int t = idx[j]; idx[j] = idx[k]; idx[k] = t; }
    int pL = (int)((double)ceilf((float)n * sgl0) + 0.001) - 1;
    pL = pL < 0 ? 0 : pL; pL = pL >= n ? n - 1 : pL;
    int pH = (int)((double)ceilf((float)n * sgl1) + 0.001) - 1;
    pH = pH < 0 ? 0 : pH; pH = pH >= n ? n - 1 : pH;
    int med = (int)(ceil((double)n * 0.5) + 0.001) - 1;
    med = med < 0 ? 0 : med; med = med >= n ? n - 1 : med;
    const float sigma_g = coeff * (values[idx[pH]] - values[idx[pL]]);
    const float lo = values[idx[med]] - width * sigma_g;
    const float hi = values[idx[med]] + width * sigma_g;
    int start = 0;
    while (start < med && values[idx[start]] < lo) ++start;
    *min_keep = start;
    int end = med + 1;
    while (end < n && values[idx[end]] <= hi) ++end;
    *max_keep = end - 1;
}

/* ---- S3: evaluateTrajectory, GPU-path semantics (kernels.cu:154-242) ------------------------------- */
typedef struct {
    int min_obs;
    float min_lh;
    int do_sigmag;
    float sgl_L, sgl_H, coeff;
    int x0, x1, y0, y1; /* start-pixel bounds [x0,x1) x [y0,y1) */
    int K;
} orc_params;

/* keep_mask (optional, T bytes): 1 where observation i was summed into the final lh/flux. */
ORC_API void orc_eval_trajectory(const orc_stack *s, const orc_params *p, int x, int y, float vx, float vy,
                                 float *lh_out, float *flux_out, int *obs_out, uint8_t *keep_mask) {
    const int T = s->T;
    float *psi_a = (float *)malloc(sizeof(float) * (size_t)T * 3);
    float *phi_a = psi_a + T, *lc = phi_a + T;
    int *idx = (int *)malloc(sizeof(int) * (size_t)T * 2);
    int *tix = idx + T; /* time index of each kept observation */
    float psi_sum = 0.0f, phi_sum = 0.0f;
    int seen = 0;
    if (keep_mask) memset(keep_mask, 0, (size_t)T);
    for (int i = 0; i < T; ++i) {
        const int cx = predict_index((float)x, vx, s->times[i]);
        const int cy = predict_index((float)y, vy, s->times[i]);
        float a, b;
        read_pp(s, i, cy, cx, &a, &b);
        if (valid(a) && valid(b)) {
            psi_sum += a; phi_sum += b;
            psi_a[seen] = a; phi_a[seen] = b; tix[seen] = i;
            if (keep_mask) keep_mask[i] = 1;
            ++seen;
        }
    }
    float lh = (phi_sum > 0) ? psi_sum / sqrtf(phi_sum) : -1.0f;
    float flux = (phi_sum > 0) ? psi_sum / phi_sum : -1.0f;
    if (!(seen < p->min_obs || seen == 0 || (p->do_sigmag && lh < p->min_lh)) && p->do_sigmag) {
        for (int i = 0; i < seen; ++i) lc[i] = (phi_a[i] != 0) ? psi_a[i] / phi_a[i] : 0.0f;
        int lo = 0, hi = seen - 1;
        orc_sigmag_indices(lc, seen, p->sgl_L, p->sgl_H, p->coeff, 2.0f, idx, &lo, &hi);
        if (lo < 0) lo = 0;
        if (hi >= seen) hi = seen - 1;
        float ps = 0.0f, ph = 0.0f;
        if (keep_mask) memset(keep_mask, 0, (size_t)T);
        for (int i = lo; i <= hi; ++i) { /* re-sum in ascending-lc order, :233-237 */
            ps += psi_a[idx[i]]; ph += phi_a[idx[i]];
            if (keep_mask) keep_mask[tix[idx[i]]] = 1;
        }
        lh = (ph > 0) ? ps / sqrtf(ph) : -1.0f;
        flux = (ph > 0) ? ps / ph : -1.0f;
    }
    *lh_out = lh; *flux_out = flux; *obs_out = seen; /* obs_count is NOT reduced by the clip (:195) */
    free(psi_a);
    free(idx);
}

/* Batch form of orc_eval_trajectory for the tests' record validation. */
ORC_API void orc_eval_many(const orc_stack *s, const orc_params *p, int64_t n, const int *x, const int *y,
                           const float *vx, const float *vy, float *lh, float *flux, int *obs) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < n; ++i) orc_eval_trajectory(s, p, x[i], y[i], vx[i], vy[i], &lh[i], &flux[i], &obs[i], NULL);
}

/* ---- S2: per-pixel top-K slots, GPU-path semantics (kernels.cu:252-332) ---------------------------- */
/* Outputs are K*P arrays in slot order ((y_i*sw + x_i)*K + r).  cand_idx = -1 for an empty slot. */
ORC_API void orc_search_slots(const orc_stack *s, const orc_params *p, const float *vx, const float *vy,
                              int64_t n_cand, int *ox, int *oy, float *ovx, float *ovy, float *olh, float *oflux,
                              int *oobs, int *ocand) {
    const int sw = p->x1 - p->x0, sh = p->y1 - p->y0, K = p->K;
#pragma omp parallel for collapse(2) schedule(dynamic, 8)
    for (int yi = 0; yi < sh; ++yi) {
        for (int xi = 0; xi < sw; ++xi) {
            const int x = xi + p->x0, y = yi + p->y0;
            const size_t base = ((size_t)yi * sw + xi) * K;
            for (int r = 0; r < K; ++r) { /* :293-301 */
                ox[base + r] = x; oy[base + r] = y; ovx[base + r] = 0.0f; ovy[base + r] = 0.0f;
                olh[base + r] = -FLT_MAX; oflux[base + r] = 0.0f; oobs[base + r] = 0; ocand[base + r] = -1;
            }
            for (int64_t c = 0; c < n_cand; ++c) {
                float lh, flux; int obs;
                orc_eval_trajectory(s, p, x, y, vx[c], vy[c], &lh, &flux, &obs, NULL);
                if (obs < p->min_obs || (p->do_sigmag && lh < p->min_lh)) continue; /* :318-320 */
                float cvx = vx[c], cvy = vy[c]; int cc = (int)c;
                for (int r = 0; r < K; ++r) { /* :324-330 strict >, displaced entry ripples down */
                    if (lh > olh[base + r]) {
                        float t;  int ti;
                        t = olh[base + r]; olh[base + r] = lh; lh = t;
                        t = oflux[base + r]; oflux[base + r] = flux; flux = t;
                        t = ovx[base + r]; ovx[base + r] = cvx; cvx = t;
                        t = ovy[base + r]; ovy[base + r] = cvy; cvy = t;
                        ti = oobs[base + r]; oobs[base + r] = obs; obs = ti;
                        ti = ocand[base + r]; ocand[base + r] = cc; cc = ti;
                    }
                }
            }
        }
    }
}

/* ---- N1: psi/phi curves (stack_search.cpp:14-31) uses Trajectory::get_x_index (common.h:71-79):
 *      position computed in double then narrowed to FLOAT before floor. ------------------------------ */
ORC_API void orc_psi_phi_curve(const orc_stack *s, int x, int y, float vx, float vy, float *out /* 2T */) {
    for (int i = 0; i < s->T; ++i) {
        const double t = s->times[i];
        const float xp = (float)((double)x + t * (double)vx + 0.5);
        const float yp = (float)((double)y + t * (double)vy + 0.5);
        float a, b;
        read_pp(s, i, (int)floorf(yp), (int)floorf(xp), &a, &b);
        out[i] = valid(a) ? a : 0.0f;
        out[i + s->T] = valid(b) ? b : 0.0f;
    }
}
