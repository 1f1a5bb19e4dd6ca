/*
 * TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Body of the DCNv2 CPU oracle, included twice by
 * dcn_oracle.c (REAL=double, REAL=float).  See dcn_oracle.c for provenance and pinning.
 *
 * Data layout: every array is dense NCHW (offset: B x dg*2*kh*kw x Ho x Wo with (dy,dx) interleaved
 * per tap; mask: B x dg*kh*kw x Ho x Wo; weight: Cout x Cin x kh x kw).
 */

#ifndef REAL
#error "include from dcn_oracle.c"
#endif

#define CAT2(a, b) a##b
#define CAT(a, b) CAT2(a, b)
#define FN(name) CAT(name, SUFFIX)

/* One deformed tap: the four bilinear corners (flat H*W index or -1 when that corner lies outside
 * the image) and their weights.  alive == 0 means the DCNv2 "whole sample is zero" rule fired
 * (p_y <= -1 || p_y >= H || p_x <= -1 || p_x >= W): upstream im2col guards the sample with
 * `h_im > -1 && w_im > -1 && h_im < height && w_im < width`. */
typedef struct {
  int alive;
  int idx[4];   /* (y0,x0) (y0,x1) (y1,x0) (y1,x1) */
  REAL wgt[4];
  REAL ly, lx;  /* fractional parts */
} FN(tap_t);

static void FN(tap_setup)(REAL py, REAL px, int H, int W, FN(tap_t) * t) {
  t->alive = (py > (REAL)-1 && px > (REAL)-1 && py < (REAL)H && px < (REAL)W);
  if (!t->alive) {
    for (int q = 0; q < 4; ++q) { t->idx[q] = -1; t->wgt[q] = 0; }
    t->ly = t->lx = 0;
    return;
  }
  const REAL fy = FLOOR(py), fx = FLOOR(px);
  const int y0 = (int)fy, x0 = (int)fx, y1 = y0 + 1, x1 = x0 + 1;
  const REAL ly = py - fy, lx = px - fx, hy = 1 - ly, hx = 1 - lx;
  t->ly = ly; t->lx = lx;
  t->idx[0] = (y0 >= 0 && x0 >= 0) ? y0 * W + x0 : -1;
  t->idx[1] = (y0 >= 0 && x1 <= W - 1) ? y0 * W + x1 : -1;
  t->idx[2] = (y1 <= H - 1 && x0 >= 0) ? y1 * W + x0 : -1;
  t->idx[3] = (y1 <= H - 1 && x1 <= W - 1) ? y1 * W + x1 : -1;
  t->wgt[0] = hy * hx; t->wgt[1] = hy * lx; t->wgt[2] = ly * hx; t->wgt[3] = ly * lx;
}

static inline REAL FN(corner)(const REAL* plane, int idx) { return idx >= 0 ? plane[idx] : (REAL)0; }

/* Forward: y[b,o,ho,wo] = bias[o] + sum_{c,k} W[o,c,k] * mask_k * bilinear(x[b,c], p_k). */
void FN(dcn_oracle_forward_)(const REAL* x, const REAL* offset, const REAL* mask, const REAL* weight,
                             const REAL* bias, REAL* y, int B, int Cin, int Cout, int H, int W, int kh,
                             int kw, int sh, int sw, int ph, int pw, int dh, int dw, int dg) {
  const int Ho = (H + 2 * ph - (dh * (kh - 1) + 1)) / sh + 1;
  const int Wo = (W + 2 * pw - (dw * (kw - 1) + 1)) / sw + 1;
  const int T = kh * kw, cpg = Cin / dg, Kdim = Cin * T;
  const long HWo = (long)Ho * Wo, HW = (long)H * W;
#pragma omp parallel
  {
    /* column block for one output row: col[(c*T+k)][wo], then a row-GEMM with W. */
    REAL* col = (REAL*)malloc(sizeof(REAL) * (size_t)Kdim * Wo);
#pragma omp for collapse(2) schedule(dynamic)
    for (int b = 0; b < B; ++b)
      for (int ho = 0; ho < Ho; ++ho) {
        for (int g = 0; g < dg; ++g)
          for (int k = 0; k < T; ++k) {
            const int i = k / kw, j = k % kw;
            const REAL* offy = offset + ((long)b * dg * 2 * T + (long)g * 2 * T + 2 * k) * HWo + (long)ho * Wo;
            const REAL* offx = offy + HWo;
            const REAL* mk = mask + ((long)b * dg * T + (long)g * T + k) * HWo + (long)ho * Wo;
            for (int wo = 0; wo < Wo; ++wo) {
              FN(tap_t) t;
              FN(tap_setup)((REAL)(ho * sh - ph + i * dh) + offy[wo], (REAL)(wo * sw - pw + j * dw) + offx[wo], H, W, &t);
              for (int cc = 0; cc < cpg; ++cc) {
                const int c = g * cpg + cc;
                const REAL* plane = x + ((long)b * Cin + c) * HW;
                REAL v = 0;
                if (t.alive)
                  v = t.wgt[0] * FN(corner)(plane, t.idx[0]) + t.wgt[1] * FN(corner)(plane, t.idx[1]) +
                      t.wgt[2] * FN(corner)(plane, t.idx[2]) + t.wgt[3] * FN(corner)(plane, t.idx[3]);
                col[(size_t)(c * T + k) * Wo + wo] = v * mk[wo];
              }
            }
          }
        for (int o = 0; o < Cout; ++o) {
          REAL* yrow = y + ((long)b * Cout + o) * HWo + (long)ho * Wo;
          const REAL bo = bias ? bias[o] : (REAL)0;
          for (int wo = 0; wo < Wo; ++wo) yrow[wo] = bo;
          const REAL* wrow = weight + (long)o * Kdim;
          for (int q = 0; q < Kdim; ++q) {
            const REAL wv = wrow[q];
            const REAL* crow = col + (size_t)q * Wo;
            for (int wo = 0; wo < Wo; ++wo) yrow[wo] += wv * crow[wo];
          }
        }
      }
    free(col);
  }
}

/* Backward: all five gradients (any output pointer may be NULL to skip it).
 *   gcol[c,k]       = sum_o W[o,c,k] gout[o]
 *   gmask[k]        = sum_c gcol[c,k] * v_c(p_k)
 *   goffset[2k]     = m_k sum_c gcol[c,k] * d v_c / d p_y     (floor-based forward difference)
 *   goffset[2k+1]   = m_k sum_c gcol[c,k] * d v_c / d p_x
 *   gx[c,corner]   += gcol[c,k] m_k w_corner                  (bilinear adjoint == upstream col2im)
 *   gW[o,c,k]       = sum_{b,px} gout[o] m_k v_c(p_k);  gb[o] = sum gout[o]
 * Serial over pixels inside one image so gx needs no atomics; OpenMP over (image) for gx and a
 * private gW per thread reduced at the end.  Deterministic for a fixed thread count. */
void FN(dcn_oracle_backward_)(const REAL* x, const REAL* offset, const REAL* mask, const REAL* weight,
                              const REAL* gout, REAL* gx, REAL* goffset, REAL* gmask, REAL* gw, REAL* gb,
                              int B, int Cin, int Cout, int H, int W, int kh, int kw, int sh, int sw, int ph,
                              int pw, int dh, int dw, int dg) {
  const int Ho = (H + 2 * ph - (dh * (kh - 1) + 1)) / sh + 1;
  const int Wo = (W + 2 * pw - (dw * (kw - 1) + 1)) / sw + 1;
  const int T = kh * kw, cpg = Cin / dg, Kdim = Cin * T;
  const long HWo = (long)Ho * Wo, HW = (long)H * W;
  if (gx) memset(gx, 0, sizeof(REAL) * (size_t)B * Cin * HW);
  if (gw) memset(gw, 0, sizeof(REAL) * (size_t)Cout * Kdim);
  if (gb) memset(gb, 0, sizeof(REAL) * (size_t)Cout);
#pragma omp parallel
  {
    REAL* gcol = (REAL*)malloc(sizeof(REAL) * (size_t)Kdim);
    REAL* colv = (REAL*)malloc(sizeof(REAL) * (size_t)Kdim); /* m_k * v_c(p_k), for gW */
    REAL* gw_private = gw ? (REAL*)calloc((size_t)Cout * Kdim, sizeof(REAL)) : NULL;
    REAL* gb_private = gb ? (REAL*)calloc((size_t)Cout, sizeof(REAL)) : NULL;
#pragma omp for schedule(static)
    for (int b = 0; b < B; ++b)
      for (int ho = 0; ho < Ho; ++ho)
        for (int wo = 0; wo < Wo; ++wo) {
          const long pix = (long)ho * Wo + wo;
          /* dgrad: gcol = W^T gout */
          for (int q = 0; q < Kdim; ++q) gcol[q] = 0;
          for (int o = 0; o < Cout; ++o) {
            const REAL go = gout[((long)b * Cout + o) * HWo + pix];
            const REAL* wrow = weight + (long)o * Kdim;
            for (int q = 0; q < Kdim; ++q) gcol[q] += wrow[q] * go;
            if (gb_private) gb_private[o] += go;
          }
          for (int g = 0; g < dg; ++g)
            for (int k = 0; k < T; ++k) {
              const int i = k / kw, j = k % kw;
              const long och = (long)b * dg * 2 * T + (long)g * 2 * T + 2 * k;
              const long mch = (long)b * dg * T + (long)g * T + k;
              const REAL m = mask[mch * HWo + pix];
              FN(tap_t) t;
              FN(tap_setup)((REAL)(ho * sh - ph + i * dh) + offset[och * HWo + pix],
                            (REAL)(wo * sw - pw + j * dw) + offset[(och + 1) * HWo + pix], H, W, &t);
              REAL acc_m = 0, acc_y = 0, acc_x = 0;
              for (int cc = 0; cc < cpg; ++cc) {
                const int c = g * cpg + cc;
                const REAL* plane = x + ((long)b * Cin + c) * HW;
                const REAL gc = gcol[c * T + k];
                REAL v = 0;
                if (t.alive) {
                  const REAL v00 = FN(corner)(plane, t.idx[0]), v01 = FN(corner)(plane, t.idx[1]);
                  const REAL v10 = FN(corner)(plane, t.idx[2]), v11 = FN(corner)(plane, t.idx[3]);
                  v = t.wgt[0] * v00 + t.wgt[1] * v01 + t.wgt[2] * v10 + t.wgt[3] * v11;
                  acc_m += gc * v;
                  acc_y += gc * ((1 - t.lx) * (v10 - v00) + t.lx * (v11 - v01));
                  acc_x += gc * ((1 - t.ly) * (v01 - v00) + t.ly * (v11 - v10));
                  if (gx) {
                    REAL* gplane = gx + ((long)b * Cin + c) * HW;
                    const REAL s = gc * m;
                    for (int q = 0; q < 4; ++q)
                      if (t.idx[q] >= 0) gplane[t.idx[q]] += s * t.wgt[q];
                  }
                }
                colv[c * T + k] = m * v;
              }
              if (gmask) gmask[mch * HWo + pix] = acc_m;
              if (goffset) {
                goffset[och * HWo + pix] = m * acc_y;
                goffset[(och + 1) * HWo + pix] = m * acc_x;
              }
            }
          if (gw_private)
            for (int o = 0; o < Cout; ++o) {
              const REAL go = gout[((long)b * Cout + o) * HWo + pix];
              REAL* grow = gw_private + (long)o * Kdim;
              for (int q = 0; q < Kdim; ++q) grow[q] += go * colv[q];
            }
        }
#pragma omp critical
    {
      if (gw_private)
        for (long q = 0; q < (long)Cout * Kdim; ++q) gw[q] += gw_private[q];
      if (gb_private)
        for (int o = 0; o < Cout; ++o) gb[o] += gb_private[o];
    }
    free(gcol); free(colv); free(gw_private); free(gb_private);
  }
}

#undef FN
#undef CAT
#undef CAT2
