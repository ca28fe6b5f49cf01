/*
 * oracle/fe_ref.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  See fe_ref.h.
 *
 * Restates OpenCV 4.x (pinned by cv2 4.13.0 in tests/test_oracle_fe.py):
 *   modules/imgproc/src/pyramids.cpp        pyrDown, 8U, 5x5 Gaussian, REFLECT_101
 *   modules/video/src/lkpyramid.cpp         calcScharrDeriv, LKTrackerInvoker,
 *                                           buildOpticalFlowPyramid, calcOpticalFlowPyrLK
 *   modules/imgproc/src/corner.cpp          cornerMinEigenVal (Sobel 3, box 3)
 *   modules/imgproc/src/featureselect.cpp   goodFeaturesToTrack
 * as called from src/processing/FeatureTracker.cpp:222-223 and :238-240.
 */
#include "fe_ref.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* cv::borderInterpolate(p, len, BORDER_REFLECT_101) */
static inline int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * len - 2 - p;
    }
    return p;
}
static inline int cv_floor(float v) { return (int)floorf(v); }
/* cvRound: round half to even (SSE cvtss2si / lrint in the default rounding mode) */
static inline int cv_round(double v) { return (int)lrint(v); }

/* -------------------------------------------------------------------- pyrDown */
void vf_pyr_down(const uint8_t* src, int w, int h, size_t pitch, uint8_t* dst,
                 size_t dpitch) {
    const int dw = (w + 1) / 2, dh = (h + 1) / 2;
    int* rows = (int*)malloc(sizeof(int) * (size_t)dw * 5);
    for (int y = 0; y < dh; ++y) {
        for (int k = 0; k < 5; ++k) { /* horizontal [1 4 6 4 1] on source row 2y-2+k */
            const uint8_t* s = src + (size_t)reflect101(2 * y - 2 + k, h) * pitch;
            int* r = rows + (size_t)k * dw;
            for (int x = 0; x < dw; ++x) {
                const int x0 = reflect101(2 * x - 2, w), x1 = reflect101(2 * x - 1, w);
                const int x2 = 2 * x, x3 = reflect101(2 * x + 1, w), x4 = reflect101(2 * x + 2, w);
                r[x] = s[x0] + 4 * s[x1] + 6 * s[x2] + 4 * s[x3] + s[x4];
            }
        }
        uint8_t* d = dst + (size_t)y * dpitch;
        for (int x = 0; x < dw; ++x) {
            const int v = rows[x] + 4 * rows[dw + x] + 6 * rows[2 * dw + x] +
                          4 * rows[3 * dw + x] + rows[4 * dw + x];
            d[x] = (uint8_t)((v + 128) >> 8);
        }
    }
    free(rows);
}

/* ----------------------------------------------------------- calcScharrDeriv */
void vf_scharr(const uint8_t* src, int w, int h, size_t pitch, int16_t* dxdy) {
    for (int y = 0; y < h; ++y) {
        const uint8_t* r0 = src + (size_t)reflect101(y - 1, h) * pitch;
        const uint8_t* r1 = src + (size_t)y * pitch;
        const uint8_t* r2 = src + (size_t)reflect101(y + 1, h) * pitch;
        for (int x = 0; x < w; ++x) {
            const int xm = reflect101(x - 1, w), xp = reflect101(x + 1, w);
            /* trow0 = 3*(r0+r2) + 10*r1 ; trow1 = r2 - r0 */
            const int t0m = (r0[xm] + r2[xm]) * 3 + r1[xm] * 10;
            const int t0p = (r0[xp] + r2[xp]) * 3 + r1[xp] * 10;
            const int t1m = r2[xm] - r0[xm], t1c = r2[x] - r0[x], t1p = r2[xp] - r0[xp];
            dxdy[((size_t)y * w + x) * 2 + 0] = (int16_t)(t0p - t0m);
            dxdy[((size_t)y * w + x) * 2 + 1] = (int16_t)((t1p + t1m) * 3 + t1c * 10);
        }
    }
}

/* ---------------------------------------------------- pyramid with aprons (LK) */
typedef struct {
    int w, h, b;       /* level size, apron width */
    uint8_t* img;      /* (h+2b) x (w+2b), REFLECT_101 apron */
    int16_t* deriv;    /* (h+2b) x (w+2b) x 2, zero apron (prev pyramid only) */
} vf_level;

static void level_make_apron(vf_level* L, const uint8_t* core, size_t pitch) {
    const int w = L->w, h = L->h, b = L->b, W = w + 2 * b;
    L->img = (uint8_t*)malloc((size_t)W * (h + 2 * b));
    for (int y = -b; y < h + b; ++y) {
        const uint8_t* s = core + (size_t)reflect101(y, h) * pitch;
        uint8_t* d = L->img + (size_t)(y + b) * W;
        for (int x = -b; x < w + b; ++x) d[x + b] = s[reflect101(x, w)];
    }
}
static void level_make_deriv(vf_level* L, const uint8_t* core, size_t pitch) {
    const int w = L->w, h = L->h, b = L->b, W = w + 2 * b;
    int16_t* tmp = (int16_t*)malloc(sizeof(int16_t) * 2 * (size_t)w * h);
    vf_scharr(core, w, h, pitch, tmp);
    L->deriv = (int16_t*)calloc((size_t)W * (h + 2 * b) * 2, sizeof(int16_t));
    for (int y = 0; y < h; ++y)
        memcpy(L->deriv + ((size_t)(y + b) * W + b) * 2, tmp + (size_t)y * w * 2,
               sizeof(int16_t) * 2 * (size_t)w);
    free(tmp);
}

#define W_BITS 14
#define DESCALE(x, n) (((x) + (1 << ((n)-1))) >> (n))

/* LKTrackerInvoker::operator() for one level, all points (lkpyramid.cpp) */
static void lk_level(const vf_level* I, const vf_level* J, int level, int max_level, int win,
                     int max_iter, double eps2, double min_eig_thr, const float* prev_xy,
                     int n, float* next_xy, uint8_t* status, float* err) {
    const float half = (win - 1) * 0.5f;
    const int b = I->b, SW = I->w + 2 * b; /* padded stride (pixels) */
    const int JW = J->w + 2 * b;
    const float FLT_SCALE = 1.f / (1 << 20);
    int16_t* Iw = (int16_t*)malloc(sizeof(int16_t) * (size_t)win * win);
    int16_t* dIw = (int16_t*)malloc(sizeof(int16_t) * 2 * (size_t)win * win);
    for (int p = 0; p < n; ++p) {
        float px = prev_xy[2 * p] * (float)(1. / (1 << level));
        float py = prev_xy[2 * p + 1] * (float)(1. / (1 << level));
        float nx, ny;
        if (level == max_level) { nx = px; ny = py; }
        else { nx = next_xy[2 * p] * 2.f; ny = next_xy[2 * p + 1] * 2.f; }
        next_xy[2 * p] = nx; next_xy[2 * p + 1] = ny;
        px -= half; py -= half;
        int ipx = cv_floor(px), ipy = cv_floor(py);
        if (ipx < -win || ipx >= I->w || ipy < -win || ipy >= I->h) {
            if (level == 0) { status[p] = 0; if (err) err[p] = 0; }
            continue;
        }
        float a = px - ipx, bb = py - ipy;
        int iw00 = cv_round((1.f - a) * (1.f - bb) * (1 << W_BITS));
        int iw01 = cv_round(a * (1.f - bb) * (1 << W_BITS));
        int iw10 = cv_round((1.f - a) * bb * (1 << W_BITS));
        int iw11 = (1 << W_BITS) - iw00 - iw01 - iw10;
        float iA11 = 0, iA12 = 0, iA22 = 0; /* acctype = float on non-NEON builds */
        for (int y = 0; y < win; ++y) {
            const uint8_t* src = I->img + (size_t)(y + ipy + b) * SW + (ipx + b);
            const int16_t* dsrc = I->deriv + ((size_t)(y + ipy + b) * SW + (ipx + b)) * 2;
            for (int x = 0; x < win; ++x, dsrc += 2) {
                int ival = DESCALE(src[x] * iw00 + src[x + 1] * iw01 + src[x + SW] * iw10 +
                                   src[x + SW + 1] * iw11, W_BITS - 5);
                int ixval = DESCALE(dsrc[0] * iw00 + dsrc[2] * iw01 + dsrc[SW * 2] * iw10 +
                                    dsrc[SW * 2 + 2] * iw11, W_BITS);
                int iyval = DESCALE(dsrc[1] * iw00 + dsrc[3] * iw01 + dsrc[SW * 2 + 1] * iw10 +
                                    dsrc[SW * 2 + 3] * iw11, W_BITS);
                Iw[y * win + x] = (int16_t)ival;
                dIw[(y * win + x) * 2] = (int16_t)ixval;
                dIw[(y * win + x) * 2 + 1] = (int16_t)iyval;
                iA11 += (float)(ixval * ixval);
                iA12 += (float)(ixval * iyval);
                iA22 += (float)(iyval * iyval);
            }
        }
        float A11 = iA11 * FLT_SCALE, A12 = iA12 * FLT_SCALE, A22 = iA22 * FLT_SCALE;
        float D = A11 * A22 - A12 * A12;
        float minEig = (A22 + A11 - sqrtf((A11 - A22) * (A11 - A22) + 4.f * A12 * A12)) /
                       (float)(2 * win * win);
        if (minEig < min_eig_thr || D < FLT_EPSILON) {
            if (level == 0) status[p] = 0;
            continue;
        }
        D = 1.f / D;
        nx -= half; ny -= half;
        float pdx = 0, pdy = 0;
        for (int j = 0; j < max_iter; ++j) {
            int inx = cv_floor(nx), iny = cv_floor(ny);
            if (inx < -win || inx >= J->w || iny < -win || iny >= J->h) {
                if (level == 0) status[p] = 0;
                break;
            }
            a = nx - inx; bb = ny - iny;
            iw00 = cv_round((1.f - a) * (1.f - bb) * (1 << W_BITS));
            iw01 = cv_round(a * (1.f - bb) * (1 << W_BITS));
            iw10 = cv_round((1.f - a) * bb * (1 << W_BITS));
            iw11 = (1 << W_BITS) - iw00 - iw01 - iw10;
            float ib1 = 0, ib2 = 0;
            for (int y = 0; y < win; ++y) {
                const uint8_t* Jp = J->img + (size_t)(y + iny + b) * JW + (inx + b);
                for (int x = 0; x < win; ++x) {
                    int diff = DESCALE(Jp[x] * iw00 + Jp[x + 1] * iw01 + Jp[x + JW] * iw10 +
                                       Jp[x + JW + 1] * iw11, W_BITS - 5) - Iw[y * win + x];
                    ib1 += (float)(diff * dIw[(y * win + x) * 2]);
                    ib2 += (float)(diff * dIw[(y * win + x) * 2 + 1]);
                }
            }
            float b1 = ib1 * FLT_SCALE, b2 = ib2 * FLT_SCALE;
            float dx = (float)((A12 * b2 - A22 * b1) * D);
            float dy = (float)((A12 * b1 - A11 * b2) * D);
            nx += dx; ny += dy;
            next_xy[2 * p] = nx + half; next_xy[2 * p + 1] = ny + half;
            if ((double)dx * dx + (double)dy * dy <= eps2) break;
            if (j > 0 && fabs(dx + pdx) < 0.01 && fabs(dy + pdy) < 0.01) {
                next_xy[2 * p] -= dx * 0.5f; next_xy[2 * p + 1] -= dy * 0.5f;
                break;
            }
            pdx = dx; pdy = dy;
        }
        if (status[p] && err && level == 0) {
            float qx = next_xy[2 * p] - half, qy = next_xy[2 * p + 1] - half;
            int inx = cv_floor(qx), iny = cv_floor(qy);
            if (inx < -win || inx >= J->w || iny < -win || iny >= J->h) {
                status[p] = 0;
                continue;
            }
            float aa = qx - inx, ab = qy - iny;
            iw00 = cv_round((1.f - aa) * (1.f - ab) * (1 << W_BITS));
            iw01 = cv_round(aa * (1.f - ab) * (1 << W_BITS));
            iw10 = cv_round((1.f - aa) * ab * (1 << W_BITS));
            iw11 = (1 << W_BITS) - iw00 - iw01 - iw10;
            float errval = 0.f;
            for (int y = 0; y < win; ++y) {
                const uint8_t* Jp = J->img + (size_t)(y + iny + b) * JW + (inx + b);
                for (int x = 0; x < win; ++x) {
                    int diff = DESCALE(Jp[x] * iw00 + Jp[x + 1] * iw01 + Jp[x + JW] * iw10 +
                                       Jp[x + JW + 1] * iw11, W_BITS - 5) - Iw[y * win + x];
                    errval += fabsf((float)diff);
                }
            }
            err[p] = errval * 1.f / (32 * win * win);
        }
    }
    free(Iw); free(dIw);
}

int vf_lk_track(const uint8_t* prev, size_t prev_pitch, const uint8_t* cur, size_t cur_pitch,
                int W, int H, int max_level, int win, int max_iter, double eps,
                double min_eig_thr, const float* prev_xy, int n, float* cur_xy,
                uint8_t* status, float* err) {
    /* calcOpticalFlowPyrLK: criteria clamp + epsilon squared (lkpyramid.cpp) */
    if (max_iter < 0) max_iter = 0; if (max_iter > 100) max_iter = 100;
    if (eps < 0.) eps = 0.; if (eps > 10.) eps = 10.;
    const double eps2 = eps * eps;
    /* buildOpticalFlowPyramid: stop when a level would not exceed the window */
    vf_level* Ip = (vf_level*)calloc((size_t)max_level + 1, sizeof(vf_level));
    vf_level* Jp = (vf_level*)calloc((size_t)max_level + 1, sizeof(vf_level));
    uint8_t *pc = NULL, *cc = NULL; /* cores of the current level */
    size_t pp = prev_pitch, cp = cur_pitch;
    const uint8_t *ps = prev, *cs = cur;
    int w = W, h = H, levels = 0;
    for (int l = 0; l <= max_level; ++l) {
        Ip[l].w = Jp[l].w = w; Ip[l].h = Jp[l].h = h; Ip[l].b = Jp[l].b = win;
        level_make_apron(&Ip[l], ps, pp);
        level_make_deriv(&Ip[l], ps, pp);
        level_make_apron(&Jp[l], cs, cp);
        levels = l + 1;
        const int nw = (w + 1) / 2, nh = (h + 1) / 2;
        if (l == max_level || nw <= win || nh <= win) break;
        uint8_t* np_ = (uint8_t*)malloc((size_t)nw * nh);
        uint8_t* nc_ = (uint8_t*)malloc((size_t)nw * nh);
        vf_pyr_down(ps, w, h, pp, np_, (size_t)nw);
        vf_pyr_down(cs, w, h, cp, nc_, (size_t)nw);
        free(pc); free(cc);
        pc = np_; cc = nc_; ps = pc; cs = cc; pp = cp = (size_t)nw; w = nw; h = nh;
    }
    free(pc); free(cc);
    const int top = levels - 1;
    for (int i = 0; i < n; ++i) { status[i] = 1; if (err) err[i] = 0.f; }
    for (int l = top; l >= 0; --l)
        lk_level(&Ip[l], &Jp[l], l, top, win, max_iter, eps2, min_eig_thr, prev_xy, n, cur_xy,
                 status, err);
    int cnt = 0;
    for (int i = 0; i < n; ++i) cnt += status[i] != 0;
    for (int l = 0; l < levels; ++l) { free(Ip[l].img); free(Ip[l].deriv); free(Jp[l].img); }
    free(Ip); free(Jp);
    return cnt;
}

/* ------------------------------------------------------------ cornerMinEigenVal */
/* Sobel(src, CV_32F, 1,0 / 0,1, ksize 3, scale = 1/(255*4*3)) as sepFilter2D executes
 * it in the AVX2/FMA dispatched build that cv2 4.13.0 selects on FMA-capable CPUs
 * (pinned empirically, tests/test_oracle_fe.py):
 *   Dx: row [-1 0 1] (exact ints), column [1 2 1]*scale -> fma(S0+S2, k1, S1*k0)
 *   Dy: row [1 2 1]*scale -> fma(k1,c, fma(k0,b, k1*a)), column [-1 0 1] -> r2 - r0
 * then cov = (dx*dx, dx*dy, dy*dy) in float, 3x3 un-normalised boxFilter whose row sum
 * is a direct double sum and whose column sum is ColumnSum<double,float>'s sliding
 * accumulator (SUM += new; out = (float)SUM; SUM -= old) carried down the whole image,
 * and calcMinEigenVal in float without contraction.                                   */
void vf_min_eigen_val(const uint8_t* img, int w, int h, size_t pitch, float* dst) {
    const float k1 = (float)(1.0 * (1.0 / 3060.0)), k0 = (float)(2.0 * (1.0 / 3060.0));
    float* dx = (float*)malloc(sizeof(float) * (size_t)w * h);
    float* dy = (float*)malloc(sizeof(float) * (size_t)w * h);
    float* rdiff = (float*)malloc(sizeof(float) * (size_t)w * 3); /* rows y-1,y,y+1 */
    float* rsm = (float*)malloc(sizeof(float) * (size_t)w * 3);
    for (int y = 0; y < h; ++y) {
        for (int k = 0; k < 3; ++k) {
            const uint8_t* s = img + (size_t)reflect101(y - 1 + k, h) * pitch;
            for (int x = 0; x < w; ++x) {
                const float a = (float)s[reflect101(x - 1, w)], b = (float)s[x],
                            c = (float)s[reflect101(x + 1, w)];
                rdiff[k * w + x] = c - a;
                rsm[k * w + x] = fmaf(k1, c, fmaf(k0, b, k1 * a));
            }
        }
        for (int x = 0; x < w; ++x) {
            dx[(size_t)y * w + x] = fmaf(rdiff[x] + rdiff[2 * w + x], k1, rdiff[w + x] * k0);
            dy[(size_t)y * w + x] = rsm[2 * w + x] - rsm[x];
        }
    }
    free(rdiff); free(rsm);
    /* row sums of the three covariance channels, in double, for rows -1..h */
    double* R = (double*)malloc(sizeof(double) * 3 * (size_t)w * (h + 2));
    for (int yy = -1; yy <= h; ++yy) {
        const int y = reflect101(yy, h);
        double* r = R + 3 * (size_t)w * (yy + 1);
        for (int x = 0; x < w; ++x) {
            double s[3] = {0, 0, 0};
            for (int k = -1; k <= 1; ++k) {
                const int xs = reflect101(x + k, w);
                const float gx = dx[(size_t)y * w + xs], gy = dy[(size_t)y * w + xs];
                const float c0 = gx * gx, c1 = gx * gy, c2 = gy * gy;
                if (k == -1) { s[0] = (double)c0; s[1] = (double)c1; s[2] = (double)c2; }
                else { s[0] += (double)c0; s[1] += (double)c1; s[2] += (double)c2; }
            }
            r[3 * x] = s[0]; r[3 * x + 1] = s[1]; r[3 * x + 2] = s[2];
        }
    }
    double* SUM = (double*)calloc(3 * (size_t)w, sizeof(double));
    for (int k = 0; k < 2; ++k)
        for (int i = 0; i < 3 * w; ++i) SUM[i] += R[3 * (size_t)w * k + i];
    for (int y = 0; y < h; ++y) {
        const double* Sp = R + 3 * (size_t)w * (y + 2);
        const double* Sm = R + 3 * (size_t)w * y;
        for (int x = 0; x < w; ++x) {
            float cov[3];
            for (int ch = 0; ch < 3; ++ch) {
                const double s0 = SUM[3 * x + ch] + Sp[3 * x + ch];
                cov[ch] = (float)s0;
                SUM[3 * x + ch] = s0 - Sm[3 * x + ch];
            }
            const float a = cov[0] * 0.5f, b = cov[1], c = cov[2] * 0.5f;
            const float t = a - c;
            const float tt = t * t, b2 = b * b;
            dst[(size_t)y * w + x] = (a + c) - sqrtf(tt + b2);
        }
    }
    free(SUM); free(R); free(dx); free(dy);
}

typedef struct { float v; int idx; } vf_cand;
static int cand_cmp(const void* pa, const void* pb) { /* greaterThanPtr */
    const vf_cand* a = (const vf_cand*)pa; const vf_cand* b = (const vf_cand*)pb;
    if (a->v > b->v) return -1;
    if (a->v < b->v) return 1;
    return (a->idx > b->idx) ? -1 : (a->idx < b->idx ? 1 : 0);
}

int vf_good_features(const uint8_t* img, int w, int h, size_t pitch, const uint8_t* mask,
                     size_t mask_pitch, int max_corners, double quality, double min_dist,
                     float* out_xy) {
    float* eig = (float*)malloc(sizeof(float) * (size_t)w * h);
    vf_min_eigen_val(img, w, h, pitch, eig);
    /* minMaxLoc(eig, 0, &maxVal, 0, 0, mask) */
    double maxVal = 0; int have = 0;
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x)
            if (!mask || mask[(size_t)y * mask_pitch + x]) {
                const double v = eig[(size_t)y * w + x];
                if (!have || v > maxVal) { maxVal = v; have = 1; }
            }
    if (!have) maxVal = 0;
    /* threshold(eig, eig, maxVal*quality, 0, THRESH_TOZERO) */
    const float thr = (float)(maxVal * quality);
    for (size_t i = 0; i < (size_t)w * h; ++i) if (!(eig[i] > thr)) eig[i] = 0.f;
    /* dilate 3x3 + collect interior local maxima */
    vf_cand* cand = (vf_cand*)malloc(sizeof(vf_cand) * (size_t)w * h);
    size_t total = 0;
    for (int y = 1; y < h - 1; ++y)
        for (int x = 1; x < w - 1; ++x) {
            const float val = eig[(size_t)y * w + x];
            if (val == 0.f) continue;
            if (mask && !mask[(size_t)y * mask_pitch + x]) continue;
            float m = val;
            for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    const float q = eig[(size_t)(y + dy) * w + x + dx];
                    if (q > m) m = q;
                }
            if (val == m) { cand[total].v = val; cand[total].idx = y * w + x; ++total; }
        }
    int ncorners = 0;
    if (total > 0) {
        qsort(cand, total, sizeof(vf_cand), cand_cmp);
        if (min_dist >= 1) {
            const int cell = cv_round(min_dist);
            const int gw = (w + cell - 1) / cell, gh = (h + cell - 1) / cell;
            /* per-cell singly linked lists of accepted corners */
            int* head = (int*)malloc(sizeof(int) * (size_t)gw * gh);
            for (int i = 0; i < gw * gh; ++i) head[i] = -1;
            int* next = (int*)malloc(sizeof(int) * total);
            float* cx = (float*)malloc(sizeof(float) * total);
            float* cy = (float*)malloc(sizeof(float) * total);
            const float md2 = (float)(min_dist * min_dist); /* compared in float below */
            const double md2d = min_dist * min_dist;
            (void)md2;
            for (size_t i = 0; i < total; ++i) {
                const int y = cand[i].idx / w, x = cand[i].idx - y * w;
                const int xc = x / cell, yc = y / cell;
                int x1 = xc - 1, y1 = yc - 1, x2 = xc + 1, y2 = yc + 1;
                if (x1 < 0) x1 = 0; if (y1 < 0) y1 = 0;
                if (x2 > gw - 1) x2 = gw - 1; if (y2 > gh - 1) y2 = gh - 1;
                int good = 1;
                for (int yy = y1; yy <= y2 && good; ++yy)
                    for (int xx = x1; xx <= x2 && good; ++xx)
                        for (int j = head[yy * gw + xx]; j >= 0; j = next[j]) {
                            const float ddx = (float)x - cx[j], ddy = (float)y - cy[j];
                            if ((double)(ddx * ddx + ddy * ddy) < md2d) { good = 0; break; }
                        }
                if (good) {
                    cx[ncorners] = (float)x; cy[ncorners] = (float)y;
                    next[ncorners] = head[yc * gw + xc]; head[yc * gw + xc] = ncorners;
                    out_xy[2 * ncorners] = (float)x; out_xy[2 * ncorners + 1] = (float)y;
                    ++ncorners;
                    if (max_corners > 0 && ncorners == max_corners) break;
                }
            }
            free(head); free(next); free(cx); free(cy);
        } else {
            for (size_t i = 0; i < total; ++i) {
                const int y = cand[i].idx / w, x = cand[i].idx - y * w;
                out_xy[2 * ncorners] = (float)x; out_xy[2 * ncorners + 1] = (float)y;
                ++ncorners;
                if (max_corners > 0 && ncorners == max_corners) break;
            }
        }
    }
    free(cand); free(eig);
    return ncorners;
}

/* FeatureTracker.cpp:47-58 + Camera.cpp:100-118 */
void vf_base_mask(int w, int h, float polar_ratio, int boundary_margin, uint8_t* mask) {
    const int top_rows = (int)(h * polar_ratio);
    const int bottom_start = (int)(h * (1.0f - polar_ratio));
    for (int y = 0; y < h; ++y) {
        const int row_ok = !(y < top_rows) && !(y >= bottom_start);
        for (int x = 0; x < w; ++x) {
            const int col_ok = boundary_margin <= 0 ||
                               (x >= boundary_margin && x < w - boundary_margin);
            mask[(size_t)y * w + x] = (row_ok && col_ok) ? 255 : 0;
        }
    }
}

/* FeatureTracker.cpp:118-126 ; Camera.cpp:120-123, 134-139 */
void vf_post_filter(const float* xy, const uint8_t* status, int n, int w, int h,
                    float polar_thr, float margin, uint8_t* keep) {
    for (int i = 0; i < n; ++i) {
        const float x = xy[2 * i], y = xy[2 * i + 1];
        const float v_ratio = y / (float)h;
        const int polar = (v_ratio < polar_thr) || (v_ratio > (1.0f - polar_thr));
        const int nearb = (x < margin) || (x > (float)w - margin) || (y < margin) ||
                          (y > (float)h - margin);
        keep[i] = (uint8_t)(status[i] && !polar && !nearb);
    }
}

/* ------------------------------------------------------------------ occupancy mask */
static void hline0(uint8_t* mask, int w, int h, int y, int x0, int x1) {
    if (y < 0 || y >= h) return;
    if (x0 < 0) x0 = 0;
    if (x1 > w - 1) x1 = w - 1;
    for (int x = x0; x <= x1; ++x) mask[(size_t)y * w + x] = 0;
}

void vf_feature_mask(uint8_t* mask, int w, int h, const float* pts, int n, int radius) {
    for (int i = 0; i < n; ++i) {
        const int cx = (int)lrintf(pts[2 * i]), cy = (int)lrintf(pts[2 * i + 1]);   /* cvRound */
        int err = 0, dx = radius, dy = 0, plus = 1, minus = (radius << 1) - 1;
        while (dx >= dy) {
            hline0(mask, w, h, cy - dy, cx - dx, cx + dx);
            hline0(mask, w, h, cy + dy, cx - dx, cx + dx);
            hline0(mask, w, h, cy - dx, cx - dy, cx + dy);
            hline0(mask, w, h, cy + dx, cx - dy, cx + dy);
            dy++;
            err += plus;
            plus += 2;
            const int m = (err <= 0) - 1;
            err -= minus & m;
            dx += m;
            minus -= m & 2;
        }
    }
}

/* ------------------------------------------------------------------ rotation RANSAC */
#ifndef M_PI
#define M_PI 3.14159265358979323846   /* -std=c11 hides the POSIX constant */
#endif
/* Camera.cpp:22-47 */
static void pixel_to_bearing(float px, float py, int W, int H, float b[3]) {
    const float u_norm = px / (float)W, v_norm = py / (float)H;
    const float lon = (float)((double)((u_norm - 0.5f) * 2.0f) * M_PI);   /* float * double -> float */
    const float lat = (float)((double)(-(v_norm - 0.5f)) * M_PI);
    const float cos_lat = cosf(lat);
    b[0] = cos_lat * sinf(lon);
    b[1] = -sinf(lat);
    b[2] = cos_lat * cosf(lon);
    const float n2 = b[0] * b[0] + b[1] * b[1] + b[2] * b[2];          /* Eigen normalize() */
    if (n2 > 0.f) { const float nn = sqrtf(n2); b[0] /= nn; b[1] /= nn; b[2] /= nn; }
}

static float det3f(const float A[9]) {
    return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) +
           A[2] * (A[3] * A[7] - A[4] * A[6]);
}

/* U V^T of the SVD of the 3x3 M (row-major); V column 2 negated if det < 0 (FeatureTracker.cpp:345-349) */
static void kabsch_rotation(const float M[9], float R[9]) {
    float B[9], V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    memcpy(B, M, sizeof(B));
    for (int sweep = 0; sweep < 30; ++sweep) {
        float off = 0.f;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                float alpha = 0.f, beta = 0.f, gamma = 0.f;
                for (int k = 0; k < 3; ++k) {
                    alpha += B[k * 3 + p] * B[k * 3 + p];
                    beta += B[k * 3 + q] * B[k * 3 + q];
                    gamma += B[k * 3 + p] * B[k * 3 + q];
                }
                const float lim = sqrtf(alpha * beta);
                if (gamma == 0.f || fabsf(gamma) <= 1e-9f * lim) continue;
                const float rel = fabsf(gamma) / (lim > 0.f ? lim : 1.f);
                if (rel > off) off = rel;
                const float zeta = (beta - alpha) / (2.f * gamma);
                const float t = (zeta >= 0.f ? 1.f : -1.f) / (fabsf(zeta) + sqrtf(1.f + zeta * zeta));
                const float c = 1.f / sqrtf(1.f + t * t), s = c * t;
                for (int k = 0; k < 3; ++k) {
                    const float bp = B[k * 3 + p], bq = B[k * 3 + q];
                    B[k * 3 + p] = c * bp - s * bq; B[k * 3 + q] = s * bp + c * bq;
                    const float vp = V[k * 3 + p], vq = V[k * 3 + q];
                    V[k * 3 + p] = c * vp - s * vq; V[k * 3 + q] = s * vp + c * vq;
                }
            }
        if (off < 1e-7f) break;
    }
    float sv[3];
    for (int j = 0; j < 3; ++j) sv[j] = sqrtf(B[j] * B[j] + B[3 + j] * B[3 + j] + B[6 + j] * B[6 + j]);
    int ord[3] = {0, 1, 2};
    for (int a = 0; a < 2; ++a)
        for (int b = a + 1; b < 3; ++b)
            if (sv[ord[b]] > sv[ord[a]]) { const int t = ord[a]; ord[a] = ord[b]; ord[b] = t; }
    float U[9], Vs[9];
    for (int j = 0; j < 3; ++j)
        for (int k = 0; k < 3; ++k) {
            U[k * 3 + j] = sv[ord[j]] > 0.f ? B[k * 3 + ord[j]] / sv[ord[j]] : 0.f;
            Vs[k * 3 + j] = V[k * 3 + ord[j]];
        }
    if (!(sv[ord[2]] > 0.f) && sv[ord[1]] > 0.f) {
        U[2] = U[3] * U[7] - U[6] * U[4];
        U[5] = U[6] * U[1] - U[0] * U[7];
        U[8] = U[0] * U[4] - U[3] * U[1];
    }
    for (int pass = 0; pass < 2; ++pass) {
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                R[i * 3 + j] = U[i * 3] * Vs[j * 3] + U[i * 3 + 1] * Vs[j * 3 + 1] + U[i * 3 + 2] * Vs[j * 3 + 2];
        if (pass == 1 || !(det3f(R) < 0.f)) break;
        Vs[2] = -Vs[2]; Vs[5] = -Vs[5]; Vs[8] = -Vs[8];
    }
}

/* Camera.cpp:89-98 after R * b1 (FeatureTracker.cpp:366-369) */
static float rotated_angle(const float R[9], const float b1[3], const float b2[3]) {
    const float rx = R[0] * b1[0] + R[1] * b1[1] + R[2] * b1[2];
    const float ry = R[3] * b1[0] + R[4] * b1[1] + R[5] * b1[2];
    const float rz = R[6] * b1[0] + R[7] * b1[1] + R[8] * b1[2];
    float c = rx * b2[0] + ry * b2[1] + rz * b2[2];
    c = fmaxf(-1.0f, fminf(1.0f, c));
    return acosf(c);
}

int vf_ransac_rotation(const float* prev_xy, const float* cur_xy, int n, int W, int H,
                       const int* samples, int iters, float threshold_deg, uint8_t* mask,
                       int* counts, int* best_iter, float* errs) {
    if (best_iter) *best_iter = -1;
    if (n < 3) {                                                   /* :258-261 */
        for (int i = 0; i < n; ++i) mask[i] = 1;
        return n;
    }
    float* b1 = (float*)malloc(sizeof(float) * 3 * (size_t)n);
    float* b2 = (float*)malloc(sizeof(float) * 3 * (size_t)n);
    for (int i = 0; i < n; ++i) {
        pixel_to_bearing(prev_xy[2 * i], prev_xy[2 * i + 1], W, H, b1 + 3 * i);
        pixel_to_bearing(cur_xy[2 * i], cur_xy[2 * i + 1], W, H, b2 + 3 * i);
    }
    const float thr = (float)((double)threshold_deg * M_PI / 180.0);   /* :305 */
    int best = 0, bi = -1;
    float bestR[9] = {0};
    for (int it = 0; it < iters; ++it) {
        float Hm[9] = {0};
        for (int s = 0; s < 3; ++s) {                              /* :335-338: H += b2 b1^T */
            const float* a = b1 + 3 * samples[3 * it + s];
            const float* b = b2 + 3 * samples[3 * it + s];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) Hm[i * 3 + j] += b[i] * a[j];
        }
        float R[9];
        kabsch_rotation(Hm, R);
        if (fabsf(det3f(R) - 1.0f) > 0.1f) { if (counts) counts[it] = -1; continue; }   /* :300-302 */
        int cnt = 0;
        for (int i = 0; i < n; ++i) cnt += rotated_angle(R, b1 + 3 * i, b2 + 3 * i) < thr;
        if (counts) counts[it] = cnt;
        if (cnt > best) { best = cnt; bi = it; memcpy(bestR, R, sizeof(R)); }   /* :314-317 */
    }
    if (bi < 0) {
        for (int i = 0; i < n; ++i) mask[i] = 1;                   /* :321-322 */
        if (errs) for (int i = 0; i < n; ++i) errs[i] = 0.f;
    } else {
        for (int i = 0; i < n; ++i) {
            const float e = rotated_angle(bestR, b1 + 3 * i, b2 + 3 * i);
            mask[i] = e < thr;
            if (errs) errs[i] = e;
        }
    }
    if (best_iter) *best_iter = bi;
    free(b1); free(b2);
    return best;
}

/* ---- feature-list filters (FeatureTracker.cpp:404-497, Frame.cpp:108-202) ---------------------- */
static int vf_cell_index(float x, float y, float cw, float ch, int cols, int rows) {
    int col = (int)(x / cw), row = (int)(y / ch);        /* :435-436 / Frame.cpp:131-132 */
    if (col > cols - 1) col = cols - 1;
    if (row > rows - 1) row = rows - 1;
    if (col < 0) col = 0;                                 /* the reference would index out of bounds */
    if (row < 0) row = 0;
    return row * cols + col;
}

void vf_cluster_filter(const float* xy, int n, int w, int h, int cols, int rows, float std_ratio, uint8_t* keep) {
    for (int i = 0; i < n; ++i) keep[i] = 1;
    if (n < 4) return;                                    /* :411 */
    const float cw = (float)w / cols, ch = (float)h / rows;
    const float diag = sqrtf(cw * cw + ch * ch);
    const float thr = diag * std_ratio;
    const int ncell = cols * rows;
    int* cell = (int*)malloc(sizeof(int) * (size_t)n);
    uint8_t* clustered = (uint8_t*)calloc((size_t)ncell, 1);
    for (int i = 0; i < n; ++i) cell[i] = vf_cell_index(xy[2 * i], xy[2 * i + 1], cw, ch, cols, rows);
    for (int c = 0; c < ncell; ++c) {
        int cnt = 0;
        float mx = 0.f, my = 0.f;
        for (int i = 0; i < n; ++i) if (cell[i] == c) { mx += xy[2 * i]; my += xy[2 * i + 1]; ++cnt; }
        if (cnt < 4) continue;                            /* :447 */
        mx /= (float)cnt; my /= (float)cnt;
        float var = 0.f;
        for (int i = 0; i < n; ++i)
            if (cell[i] == c) { const float dx = xy[2 * i] - mx, dy = xy[2 * i + 1] - my; var += (dx * dx + dy * dy); }
        var /= (float)cnt;
        if (sqrtf(var) < thr) clustered[c] = 1;           /* :472-474 */
    }
    for (int i = 0; i < n; ++i) if (clustered[cell[i]]) keep[i] = 0;
    free(cell); free(clustered);
}

void vf_grid_cap(const float* xy, const int* track, int n, int w, int h, int cols, int rows, int max_per_cell, uint8_t* keep) {
    const float cw = (float)w / cols, ch = (float)h / rows;
    const int ncell = cols * rows;
    int* count = (int*)calloc((size_t)ncell + 1, sizeof(int));
    int* cell = (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; ++i) {
        const float x = xy[2 * i], y = xy[2 * i + 1];
        keep[i] = 0;
        cell[i] = -1;
        if (x < 0 || x >= w || y < 0 || y >= h) continue;  /* Frame.cpp:125-128 */
        cell[i] = vf_cell_index(x, y, cw, ch, cols, rows);
        count[cell[i] + 1]++;
    }
    for (int c = 0; c < ncell; ++c) count[c + 1] += count[c];
    int* list = (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    int* fill = (int*)malloc(sizeof(int) * (size_t)ncell);
    for (int c = 0; c < ncell; ++c) fill[c] = count[c];
    for (int i = 0; i < n; ++i) if (cell[i] >= 0) list[fill[cell[i]]++] = i;   /* push_back in feature order */
    for (int c = 0; c < ncell; ++c) {
        int* v = list + count[c];
        int sz = count[c + 1] - count[c];
        if (sz > max_per_cell) {
            /* std::sort(cell, comp = track[a] > track[b]) as libstdc++'s __insertion_sort */
            for (int i = 1; i < sz; ++i) {
                const int val = v[i];
                if (track[val] > track[v[0]]) {
                    for (int k = i; k > 0; --k) v[k] = v[k - 1];
                    v[0] = val;
                } else {
                    int k = i;
                    while (track[val] > track[v[k - 1]]) { v[k] = v[k - 1]; --k; }
                    v[k] = val;
                }
            }
            sz = max_per_cell;                              /* cell.resize */
        }
        for (int k = 0; k < sz; ++k) keep[v[k]] = 1;
    }
    free(count); free(cell); free(list); free(fill);
}
