/*
 * TEST INFRASTRUCTURE ONLY — a plain-C CPU restatement of the reference's gradient-sync + update
 * algorithms.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and there only as the checker / CPU baseline.  The product path
 * (mxnet_maintenance_b200 + libb200kv.so) never links, imports or executes it.
 *
 * Parity pin: every function below is compared BIT-FOR-BIT with the reference's own functors
 * compiled from /root/reference (oracle/_ref/libmxref.so, see oracle/ref_shim.cc) in
 * tests/test_oracle_vs_ref.py, and against committed golden vectors generated from those functors
 * (tests/golden/, generator tests/golden/make_golden.py) plus the closed-form known answers the
 * reference's own tests hold (tests/test_host_logic.py, tests/test_kvstore_api_gpu.py).
 *
 * All arithmetic is IEEE fp32 with no contraction (-ffp-contract=off) in the reference's evaluation
 * order; fp16 goes through F16C round-to-nearest-even exactly as mshadow::half::half_t does
 * (3rdparty/mshadow/mshadow/half.h:35-39,64-85,340-349).
 */
#include <immintrin.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>

typedef uint16_t f16;
static inline float h2f(f16 h) { return _cvtsh_ss(h); }
static inline f16 f2h(float f) { return _cvtss_sh(f, 0); }
/* half_t binary add: float add, round to half (half.h:64-85) */
static inline f16 hadd(f16 a, f16 b) { return f2h(h2f(a) + h2f(b)); }

/* mshadow_op::clip::Map, src/operator/mshadow_op.h:1022-1032 */
static inline float clipf(float x, float bound) {
  if (x > bound) return bound;
  if (x < -bound) return -bound;
  return x;
}

/* ------------------------------------------------------------------------------------------
 * Dense reduce: CommCPU::ReduceSumCPU / ReduceSumCPUImpl, src/kvstore/comm.h:357-410.
 * in_0 += in_1 (+ in_2 (+ in_3 (+ in_4))) in groups of up to four; the right-hand side is summed
 * left to right first, then added to in_0.
 * ------------------------------------------------------------------------------------------ */
static void reduce_f32(float** d, int n, size_t off, size_t size) {
  float* a = d[0] + off;
  for (int i = 1; i < n; i += 4) {
    int rem = n - i;
    const float *b = d[i] + off, *c = rem > 1 ? d[i + 1] + off : 0, *e = rem > 2 ? d[i + 2] + off : 0,
                *f = rem > 3 ? d[i + 3] + off : 0;
    if (rem == 1) for (size_t k = 0; k < size; ++k) a[k] += b[k];
    else if (rem == 2) for (size_t k = 0; k < size; ++k) a[k] += b[k] + c[k];
    else if (rem == 3) for (size_t k = 0; k < size; ++k) a[k] += (b[k] + c[k]) + e[k];
    else for (size_t k = 0; k < size; ++k) a[k] += ((b[k] + c[k]) + e[k]) + f[k];
  }
}
static void reduce_f16(f16** d, int n, size_t off, size_t size) {
  f16* a = d[0] + off;
  for (int i = 1; i < n; i += 4) {
    int rem = n - i;
    const f16 *b = d[i] + off, *c = rem > 1 ? d[i + 1] + off : 0, *e = rem > 2 ? d[i + 2] + off : 0,
              *f = rem > 3 ? d[i + 3] + off : 0;
    if (rem == 1) for (size_t k = 0; k < size; ++k) a[k] = hadd(a[k], b[k]);
    else if (rem == 2) for (size_t k = 0; k < size; ++k) a[k] = hadd(a[k], hadd(b[k], c[k]));
    else if (rem == 3) for (size_t k = 0; k < size; ++k) a[k] = hadd(a[k], hadd(hadd(b[k], c[k]), e[k]));
    else for (size_t k = 0; k < size; ++k) a[k] = hadd(a[k], hadd(hadd(hadd(b[k], c[k]), e[k]), f[k]));
  }
}

int mxref_reduce_sum(int dtype, int n, void** ptrs, int64_t size, int64_t bigarray_bound,
                     int nthread) {
  if (dtype != 0 && dtype != 2) return -1;
  size_t total = (size_t)size;
  size_t step = (size_t)bigarray_bound < (size_t)(4 << 10) ? (size_t)bigarray_bound : (size_t)(4 << 10);
  if (total < (size_t)bigarray_bound || nthread <= 1) {
    if (dtype == 0) reduce_f32((float**)ptrs, n, 0, total); else reduce_f16((f16**)ptrs, n, 0, total);
    return 0;
  }
  long ntask = (long)((total + step - 1) / step);
#pragma omp parallel for schedule(static) num_threads(nthread)
  for (long j = 0; j < ntask; ++j) {
    size_t begin = (size_t)j * step; if (begin > total) begin = total;
    size_t end = ((size_t)j + 1) * step; if (end > total) end = total;
    if (dtype == 0) reduce_f32((float**)ptrs, n, begin, end - begin);
    else reduce_f16((f16**)ptrs, n, begin, end - begin);
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * MultiSGDKernel::Map, src/operator/optimizer_op-inl.h:225-258.
 *   mom = momentum*mom - lr*wd*w - lr*clip(rescale*g, c)        (c >= 0)
 *   mom = momentum*mom - lr*wd*w - lr*rescale*g                 (c <  0)
 *   w += mom;  has_mp: w lives in fp32 weights32, out (fp16) = half(w)
 * dtype 0: fp32 everywhere.  dtype 2 with has_mp: fp16 weights/grads, fp32 mom + master.
 * ------------------------------------------------------------------------------------------ */
static inline float sgd_elem(float w, float* mom_io, float g, float lr, float wd, float momentum,
                             float rescale, float clip) {
  float mom = *mom_io;
  if (clip >= 0.0f) mom = momentum * mom - lr * wd * w - lr * clipf(rescale * g, clip);
  else mom = momentum * mom - lr * wd * w - lr * rescale * g;
  *mom_io = mom;
  return w + mom;
}

int mxref_multi_sgd(int dtype, int has_mom, int has_mp, int count, void** weights, void** grads,
                    void** moms, void** w32, void** outs, const int64_t* sizes, const float* lrs,
                    const float* wds, float momentum, float rescale, float clip, int nthread) {
  if (!((dtype == 0 && !has_mp) || (dtype == 2 && has_mp))) return -1;
  for (int t = 0; t < count; ++t) {
    const int64_t n = sizes[t];
    const float lr = lrs[t], wd = wds[t];
    float* mom = has_mom ? (float*)moms[t] : 0;
    if (dtype == 0) {
      const float* w = (const float*)weights[t]; const float* g = (const float*)grads[t];
      float* out = (float*)outs[t];
#pragma omp parallel for num_threads(nthread)
      for (int64_t i = 0; i < n; ++i) {
        float m = has_mom ? mom[i] : 0.0f;
        float nw = sgd_elem(w[i], &m, g[i], lr, wd, momentum, rescale, clip);
        if (has_mom) mom[i] = m;
        out[i] = nw;
      }
    } else {
      const f16* g = (const f16*)grads[t]; float* master = (float*)w32[t]; f16* out = (f16*)outs[t];
#pragma omp parallel for num_threads(nthread)
      for (int64_t i = 0; i < n; ++i) {
        float m = has_mom ? mom[i] : 0.0f;
        float nw = sgd_elem(master[i], &m, h2f(g[i]), lr, wd, momentum, rescale, clip);
        if (has_mom) mom[i] = m;
        master[i] = nw;
        out[i] = f2h(nw);
      }
    }
  }
  return 0;
}

/* SGDMomKernel::Map, optimizer_op-inl.h:601-620 (single tensor, fp32) */
int mxref_sgd_mom(int64_t n, float* out, float* mom, const float* w, const float* g, float clip,
                  float momentum, float lr, float wd, float rescale) {
  for (int64_t i = 0; i < n; ++i) {
    if (clip >= 0.0f) mom[i] = momentum * mom[i] - lr * wd * w[i] - lr * clipf(rescale * g[i], clip);
    else mom[i] = momentum * mom[i] - lr * wd * w[i] - lr * rescale * g[i];
    out[i] = w[i] + mom[i];
  }
  return 0;
}

/* MP_SGDMomKernel::Map, optimizer_op-inl.h:700-724 */
int mxref_mp_sgd_mom(int64_t n, void* out16, float* mom, const void* w16, const void* g16,
                     float* w32, float clip, float momentum, float lr, float wd, float rescale) {
  (void)w16;
  const f16* g = (const f16*)g16; f16* out = (f16*)out16;
  for (int64_t i = 0; i < n; ++i) {
    float m = mom[i];
    float nw = sgd_elem(w32[i], &m, h2f(g[i]), lr, wd, momentum, rescale, clip);
    mom[i] = m; w32[i] = nw; out[i] = f2h(nw);
  }
  return 0;
}

/* AdamUpdateKernel::Map, optimizer_op-inl.h:1293-1315 (fp32).
 * Bias correction is folded into lr by the Python caller (optimizer.py:1616-1619). */
int mxref_adam(int64_t n, float* out, float* mean, float* var, const float* w, const float* g,
               float clip, float rescale, float beta1, float beta2, float lr, float wd, float eps,
               int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < n; ++i) {
    float gr = g[i] * rescale + w[i] * wd;
    if (clip >= 0.f) gr = clipf(gr, clip);
    mean[i] = beta1 * mean[i] + (1.f - beta1) * gr;
    var[i] = beta2 * var[i] + (1.f - beta2) * gr * gr;
    out[i] = w[i] - lr * mean[i] / (sqrtf(var[i]) + eps);
  }
  return 0;
}

/* _multi_adamw_update / _multi_mp_adamw_update: MultiMPAdamWKernel::Map,
 * src/operator/contrib/adamw-inl.h:346-377.
 *   sg = rescale*g; clip; mean = b1*(m - sg) + sg; adj = sg*sg; var = b2*(v - adj) + adj;
 *   w = w - eta*(lr*mean/(sqrt(var)+eps) + wd*w)
 * A non-finite or zero rescale_grad skips the update (MPUpdate, adamw-inl.h:447-460). */
int mxref_multi_adamw(int dtype, int has_mp, int count, void** weights, void** grads, void** means,
                      void** vars, void** w32, void** outs, const int64_t* sizes, const float* lrs,
                      const float* wds, const float* etas, float beta1, float beta2, float eps,
                      float rescale, float clip, int nthread) {
  if (!((dtype == 0 && !has_mp) || (dtype == 2 && has_mp))) return -1;
  if (!isfinite(rescale) || rescale == 0) return 0;
  for (int t = 0; t < count; ++t) {
    const int64_t n = sizes[t];
    float* mean = (float*)means[t]; float* var = (float*)vars[t];
    const float lr = lrs[t], wd = wds[t], eta = etas[t];
#pragma omp parallel for num_threads(nthread)
    for (int64_t i = 0; i < n; ++i) {
      float w = has_mp ? ((float*)w32[t])[i] : ((const float*)weights[t])[i];
      float sg = rescale * (dtype == 0 ? ((const float*)grads[t])[i] : h2f(((const f16*)grads[t])[i]));
      if (clip >= 0.0f) sg = clipf(sg, clip);
      const float m = beta1 * (mean[i] - sg) + sg;
      const float adj = sg * sg;
      const float v = beta2 * (var[i] - adj) + adj;
      mean[i] = m; var[i] = v;
      w = w - eta * (lr * m / (sqrtf(v) + eps) + wd * w);
      if (has_mp) { ((float*)w32[t])[i] = w; ((f16*)outs[t])[i] = f2h(w); }
      else ((float*)outs[t])[i] = w;
    }
  }
  return 0;
}

/* multi_lars: MultiLARSKernel::Map, src/operator/contrib/multi_lars-inl.h:60-72 */
int mxref_multi_lars(int n, float* out, const float* lrs, const float* w2, const float* g2,
                     const float* wds, float eta, float eps, float rescale) {
  for (int i = 0; i < n; ++i) {
    float w_norm = sqrtf(w2[i]);
    int valid = w_norm > 0. && g2[i] > 0.;
    out[i] = valid ? lrs[i] * eta * w_norm / (sqrtf(g2[i]) * rescale + wds[i] * w_norm + eps) : lrs[i];
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * multi_sum_sq CPU: CalcSumSq, src/operator/contrib/multi_sum_sq.cc:68-84 — one thread per array,
 * strictly sequential fp32 accumulation inside an array.
 * ------------------------------------------------------------------------------------------ */
int mxref_multi_sum_sq(int dtype, int count, void** arrays, const int64_t* sizes, float* out,
                       int nthread) {
  if (dtype != 0 && dtype != 2) return -1;
#pragma omp parallel for num_threads(nthread)
  for (int t = 0; t < count; ++t) {
    float sum = 0;
    if (dtype == 0) {
      const float* a = (const float*)arrays[t];
      for (int64_t j = 0; j < sizes[t]; ++j) sum += a[j] * a[j];
    } else {
      /* half_t * half_t rounds the product to half first (half.h operator*), then += in float */
      const f16* a = (const f16*)arrays[t];
      for (int64_t j = 0; j < sizes[t]; ++j) sum += h2f(f2h(h2f(a[j]) * h2f(a[j])));
    }
    out[t] = sum;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * multi_lamb_update / multi_mp_lamb_update CPU: MultiLAMB, multi_lamb-inl.h:270-340 driving
 * MultiLAMBKernelStep1/2, multi_lamb.cc:33-118.
 *   r1_t = sum(weights_t^2)  (of the op's `weights` input: fp16 in MP mode)
 *   step1: g = rescale*grad; clip; m,v update; u = m_hat/(sqrt(v_hat)+eps) + wd*w  -> temp_g
 *   r2_t = sum(u_t^2)
 *   step2: ratio = (r1==0||r2==0) ? 1 : sqrt(r1)/sqrt(r2) (r1 clamped to [lower,upper] when >= 0);
 *          w -= lr*ratio*u
 * ------------------------------------------------------------------------------------------ */
int mxref_multi_lamb(int dtype, int has_mp, int count, void** weights, void** grads, void** means,
                     void** vars, void** w32, void** outs, const int64_t* sizes, const float* lrs,
                     const float* wds, const int* steps, float beta1, float beta2, float eps,
                     float rescale, float clip, float lower, float upper, int bias_correction,
                     int nthread, float* r1_out, float* r2_out) {
  if (!((dtype == 0 && !has_mp) || (dtype == 2 && has_mp))) return -1;
  float* r1 = (float*)malloc(sizeof(float) * (size_t)count);
  float* r2 = (float*)malloc(sizeof(float) * (size_t)count);
  float** tmp = (float**)malloc(sizeof(float*) * (size_t)count);
  mxref_multi_sum_sq(dtype, count, weights, sizes, r1, nthread);
  for (int t = 0; t < count; ++t) {
    const int64_t n = sizes[t];
    float* u = tmp[t] = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
    float* mean = (float*)means[t]; float* var = (float*)vars[t];
    const float wd = wds[t];
    const float stepf = (float)steps[t];
#pragma omp parallel for num_threads(nthread)
    for (int64_t i = 0; i < n; ++i) {
      float w = has_mp ? ((float*)w32[t])[i] : ((const float*)weights[t])[i];
      float sg = (dtype == 0 ? ((const float*)grads[t])[i] : h2f(((const f16*)grads[t])[i])) * rescale;
      if (clip >= 0.0f) sg = clipf(sg, clip);
      float m = beta1 * mean[i] + (1.0f - beta1) * sg;
      float v = beta2 * var[i] + (1.0f - beta2) * sg * sg;
      mean[i] = m; var[i] = v;
      float g;
      if (bias_correction) {
        float mh = m / (1.0f - powf(beta1, stepf));
        float vh = v / (1.0f - powf(beta2, stepf));
        g = mh / (sqrtf(vh) + eps) + wd * w;
      } else {
        g = m / (sqrtf(v) + eps) + wd * w;
      }
      u[i] = g;
    }
  }
  mxref_multi_sum_sq(0, count, (void**)tmp, sizes, r2, nthread);
  for (int t = 0; t < count; ++t) {
    const int64_t n = sizes[t];
    float a = sqrtf(r1[t]), b = sqrtf(r2[t]);
    if (lower >= 0) a = a > lower ? a : lower;
    if (upper >= 0) a = a < upper ? a : upper;
    float r = (a == 0.0f || b == 0.0f) ? 1.0f : a / b;
    float lr_adj = lrs[t] * r;
    const float* u = tmp[t];
#pragma omp parallel for num_threads(nthread)
    for (int64_t i = 0; i < n; ++i) {
      if (has_mp) {
        float w = ((float*)w32[t])[i];
        w -= lr_adj * u[i];
        ((float*)w32[t])[i] = w;
        ((f16*)outs[t])[i] = f2h(w);
      } else {
        float w = ((const float*)weights[t])[i];
        w -= lr_adj * u[i];
        ((float*)outs[t])[i] = w;
      }
    }
    free(tmp[t]);
  }
  if (r1_out) memcpy(r1_out, r1, sizeof(float) * (size_t)count);
  if (r2_out) memcpy(r2_out, r2, sizeof(float) * (size_t)count);
  free(r1); free(r2); free(tmp);
  return 0;
}

/* SGDKernel / MP_SGDKernel::Map, optimizer_op-inl.h:382-399,656-679: single tensor, no momentum.
 * w32 != NULL: multi-precision (fp16 out, fp32 master). */
int mxref_sgd_update(int64_t n, void* out, const void* w, const void* g, float* w32, float clip, float lr,
                     float wd, float rescale) {
  if (w32) {
    const f16* g16 = (const f16*)g;
    f16* o16 = (f16*)out;
    for (int64_t i = 0; i < n; ++i) {
      float x = w32[i];
      if (clip >= 0.0f) x = (1.f - lr * wd) * x - (lr) * clipf(rescale * h2f(g16[i]), clip);
      else x = (1.f - lr * wd) * x - (lr * rescale) * h2f(g16[i]);
      w32[i] = x;
      o16[i] = f2h(x);
    }
  } else {
    const float* wf = (const float*)w;
    const float* gf = (const float*)g;
    float* of = (float*)out;
    for (int64_t i = 0; i < n; ++i) {
      if (clip >= 0.0f) of[i] = (1.f - lr * wd) * wf[i] - (lr) * clipf(rescale * gf[i], clip);
      else of[i] = (1.f - lr * wd) * wf[i] - (lr * rescale) * gf[i];
    }
  }
  return 0;
}

/* LambUpdatePhaseOneKernel / LambUpdatePhaseTwoKernel::Map, optimizer_op-inl.h:1622-1649,1706-1730 (fp32);
 * beta^t = (float)pow((double)beta, t) as LambUpdatePhaseOne computes it (:1661-1662); note the
 * double-precision division of the mean's bias correction (`1.` literal, :1643). */
int mxref_lamb_phase1(int64_t n, float* out, float* mean, float* var, const float* w, const float* g,
                      float clip, float rescale, float beta1, float beta2, float wd, float eps, int t,
                      int bias_correction) {
  const float b1t = (float)pow((double)beta1, (double)t), b2t = (float)pow((double)beta2, (double)t);
  for (int64_t i = 0; i < n; ++i) {
    float gr = g[i] * rescale;
    if (clip >= 0.f) gr = clipf(gr, clip);
    mean[i] = beta1 * mean[i] + (1.f - beta1) * gr;
    var[i] = beta2 * var[i] + (1.f - beta2) * gr * gr;
    float u = mean[i] / (sqrtf(var[i]) + eps) + wd * w[i];
    if (bias_correction) {
      float mh = (float)(mean[i] / (1. - b1t));
      float vh = var[i] / (1 - b2t);
      u = mh / (sqrtf(vh) + eps) + wd * w[i];
    }
    out[i] = u;
  }
  return 0;
}

int mxref_lamb_phase2(int64_t n, float* out, const float* w, const float* g, float r1, float r2, float lr,
                      float lower, float upper) {
  float nr1 = r1;
  if (lower >= 0) nr1 = nr1 > lower ? nr1 : lower;
  if (upper >= 0) nr1 = nr1 < upper ? nr1 : upper;
  if (nr1 == 0.0f || r2 == 0.0f) lr = lr * 1.0f;
  else lr = lr * nr1 / r2;
  for (int64_t i = 0; i < n; ++i) out[i] = w[i] - lr * g[i];
  return 0;
}

/* SGDMomDnsRspDnsKernel<req,cpu>::Map, optimizer_op-inl.h:747-774 — lazy update: only the rows
 * named by the gradient's indices are touched. */
int mxref_sgd_mom_rsp(int64_t nnr, int64_t row_len, float* out, float* mom, const float* w,
                      const int64_t* idx, const float* g, float clip, float momentum, float lr,
                      float wd, float rescale, int nthread) {
  const float rate = lr * wd;
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < nnr; ++i) {
    for (int64_t j = 0; j < row_len; ++j) {
      int64_t di = idx[i] * row_len + j, gi = i * row_len + j;
      if (clip >= 0.0f) mom[di] = momentum * mom[di] - rate * w[di] - lr * clipf(rescale * g[gi], clip);
      else mom[di] = momentum * mom[di] - rate * w[di] - lr * rescale * g[gi];
      out[di] = w[di] + mom[di];
    }
  }
  return 0;
}

/* SGDDnsRspKernel<req,cpu>::Map, optimizer_op-inl.h:452-475 — the momentum-free lazy update; note the
 * (1 - lr*wd)*w form, which rounds differently from the multi-tensor op's mom = 0 form. */
int mxref_sgd_rsp(int64_t nnr, int64_t row_len, float* out, const float* w, const int64_t* idx,
                  const float* g, float clip, float lr, float wd, float rescale, int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < nnr; ++i) {
    for (int64_t j = 0; j < row_len; ++j) {
      int64_t di = idx[i] * row_len + j, gi = i * row_len + j;
      if (clip >= 0.0f) out[di] = (1.f - lr * wd) * w[di] - (lr) * clipf(rescale * g[gi], clip);
      else out[di] = (1.f - lr * wd) * w[di] - (lr * rescale) * g[gi];
    }
  }
  return 0;
}

/* SGDMomStdDnsRspDnsKernel<req,cpu>::Map (optimizer_op.cc:108-140): standard update — EVERY row of the
 * [num_rows, row_len] weight is updated, rows absent from idx with a zero gradient. */
int mxref_sgd_mom_std_rsp(int64_t num_rows, int64_t nnr, int64_t row_len, float* out, float* mom, const float* w,
                          const int64_t* idx, const float* g, float clip, float momentum, float lr, float wd,
                          float rescale, int nthread) {
  const float rate = lr * wd;
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < num_rows; ++i) {
    int64_t lo = 0, hi = nnr - 1, p = -1;
    while (lo <= hi) { int64_t m = (lo + hi) / 2; if (idx[m] == i) { p = m; break; } if (idx[m] < i) lo = m + 1; else hi = m - 1; }
    for (int64_t j = 0; j < row_len; ++j) {
      const int64_t di = i * row_len + j;
      const float grad = p >= 0 ? g[p * row_len + j] : 0.f;
      if (clip >= 0.0f) mom[di] = momentum * mom[di] - rate * w[di] - lr * clipf(rescale * grad, clip);
      else mom[di] = momentum * mom[di] - rate * w[di] - lr * rescale * grad;
      out[di] = w[di] + mom[di];
    }
  }
  return 0;
}

/* AdamStdDnsRspDnsKernel<req,cpu>::Map (optimizer_op.cc:196-229) */
int mxref_adam_std_rsp(int64_t num_rows, int64_t nnr, int64_t row_len, float* out, float* mean, float* var,
                       const float* w, const int64_t* idx, const float* g, float clip, float beta1, float beta2,
                       float lr, float wd, float eps, float rescale, int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < num_rows; ++i) {
    int64_t lo = 0, hi = nnr - 1, p = -1;
    while (lo <= hi) { int64_t m = (lo + hi) / 2; if (idx[m] == i) { p = m; break; } if (idx[m] < i) lo = m + 1; else hi = m - 1; }
    for (int64_t j = 0; j < row_len; ++j) {
      const int64_t di = i * row_len + j;
      const float gr = p >= 0 ? (float)(g[p * row_len + j] * rescale + w[di] * wd) : (float)(w[di] * wd);
      if (clip >= 0.0f) {
        const float c = clipf(gr, clip);
        mean[di] = beta1 * mean[di] + (1.f - beta1) * c;
        var[di] = beta2 * var[di] + (1.f - beta2) * (c * c);
      } else {
        mean[di] = beta1 * mean[di] + (1.f - beta1) * gr;
        var[di] = beta2 * var[di] + (1.f - beta2) * (gr * gr);
      }
      out[di] = w[di] - lr * mean[di] / (sqrtf(var[di]) + eps);
    }
  }
  return 0;
}

/* AdamDnsRspDnsKernel<req,cpu>::Map, optimizer_op-inl.h:1352-1382 */
int mxref_adam_rsp(int64_t nnr, int64_t row_len, float* out, float* mean, float* var,
                   const float* w, const int64_t* idx, const float* g, float clip, float beta1,
                   float beta2, float lr, float wd, float eps, float rescale, int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < nnr; ++i) {
    const int64_t ro = idx[i] * row_len;
    for (int64_t j = 0; j < row_len; ++j) {
      int64_t di = ro + j, gi = i * row_len + j;
      float gr = g[gi] * rescale + w[di] * wd;
      if (clip >= 0.0f) {
        float c = clipf(gr, clip);
        mean[di] = beta1 * mean[di] + (1.f - beta1) * c;
        var[di] = beta2 * var[di] + (1.f - beta2) * (c * c);
      } else {
        mean[di] = beta1 * mean[di] + (1.f - beta1) * gr;
        var[di] = beta2 * var[di] + (1.f - beta2) * gr * gr;
      }
      out[di] = w[di] - lr * mean[di] / (sqrtf(var[di]) + eps);
    }
  }
  return 0;
}

/* AdagradDnsRspDnsKernel<cpu>::Map, optimizer_op-inl.h:2514-2539 (lazy rows, wd must be 0) */
int mxref_adagrad_rsp(int64_t nnr, int64_t row_len, float* out, float* hist, const float* w,
                      const int64_t* idx, const float* g, float clip, float eps, float lr, float rescale,
                      int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < nnr; ++i) {
    for (int64_t j = 0; j < row_len; ++j) {
      int64_t di = idx[i] * row_len + j, gi = i * row_len + j;
      float gr = g[gi] * rescale;
      if (clip >= 0.0f) gr = clipf(gr, clip);
      const float sq = gr * gr;
      hist[di] += sq;
      const float div = gr / sqrtf(hist[di] + eps);
      out[di] = w[di] - div * lr;
    }
  }
  return 0;
}

/* SparseRetainRspThreadKernel::Map, src/operator/tensor/sparse_retain-inl.h:121-150:
 * out_idx[i] = ids[i] always; row copied when present in the source, else left as the caller's zeros. */
int mxref_sparse_retain(int64_t nids, float* out_data, int64_t* out_idx, const float* in_data,
                        const int64_t* in_idx, const int64_t* ids, int64_t nnr, int64_t row_len) {
  for (int64_t i = 0; i < nids; ++i) {
    int64_t irow = ids[i];
    int64_t j = -1, left = 0, right = nnr - 1;
    while (left <= right) {
      int64_t m = left + (right - left) / 2;
      if (in_idx[m] == irow) { j = m; break; }
      else if (in_idx[m] < irow) left = m + 1;
      else right = m - 1;
    }
    out_idx[i] = ids[i];
    if (j >= 0) memcpy(out_data + i * row_len, in_data + j * row_len, sizeof(float) * (size_t)row_len);
  }
  return 0;
}

void mxref_f32_to_f16(int64_t n, const float* in, void* out) {
  f16* o = (f16*)out;
  for (int64_t i = 0; i < n; ++i) o[i] = f2h(in[i]);
}

/* ------------------------------------------------------------------------------------------
 * The functions below restate code that is bound to NDArray/Engine classes in the reference and
 * therefore cannot be lifted into oracle/_ref; they exist only in this library.
 * ------------------------------------------------------------------------------------------ */

static int cmp_i64(const void* a, const void* b) {
  int64_t x = *(const int64_t*)a, y = *(const int64_t*)b;
  return (x > y) - (x < y);
}

/* UniqueImpl<cpu>, src/kvstore/kvstore_utils.cc:32-44: sort ascending + std::unique.
 * ids is overwritten with the sorted unique values; returns their count. */
int64_t oracle_unique_i64(int64_t n, int64_t* ids) {
  if (n <= 0) return 0;
  qsort(ids, (size_t)n, sizeof(int64_t), cmp_i64);
  int64_t m = 1;
  for (int64_t i = 1; i < n; ++i) if (ids[i] != ids[m - 1]) ids[m++] = ids[i];
  return m;
}

/* ElementwiseSumRsp + GetUniqueRspRowIdx + ElementwiseSumRspImpl, src/ndarray/ndarray_function.cc:59-176.
 * ninputs row-sparse arrays (idx[k] sorted unique int64 of length nnr[k]; vals[k] nnr[k] x row_len).
 * Arrays with nnr 0 are "storage not initialised" and skipped (:81,129).  out_idx must hold
 * sum(nnr) entries, out_vals sum(nnr) x row_len; returns the number of output rows.
 * Row values: out_row = 0; for each input in order: out_row += in_row (:172, :100-102). */
int64_t oracle_rsp_sum(int ninputs, const int64_t** idx, const float** vals, const int64_t* nnr,
                       int64_t row_len, int64_t* out_idx, float* out_vals, int nthread) {
  int64_t total = 0;
  for (int k = 0; k < ninputs; ++k) { memcpy(out_idx + total, idx[k], sizeof(int64_t) * (size_t)nnr[k]); total += nnr[k]; }
  int64_t nu = oracle_unique_i64(total, out_idx);
  memset(out_vals, 0, sizeof(float) * (size_t)(nu * row_len));
#pragma omp parallel for num_threads(nthread) schedule(static)
  for (int64_t r = 0; r < nu; ++r) {
    float* o = out_vals + r * row_len;
    for (int k = 0; k < ninputs; ++k) {
      if (nnr[k] == 0) continue;
      /* lower_bound */
      int64_t lo = 0, hi = nnr[k];
      while (lo < hi) { int64_t m = lo + (hi - lo) / 2; if (idx[k][m] < out_idx[r]) lo = m + 1; else hi = m; }
      if (lo < nnr[k] && idx[k][lo] == out_idx[r]) {
        const float* in = vals[k] + lo * row_len;
        for (int64_t j = 0; j < row_len; ++j) o[j] += in[j];
      }
    }
  }
  return nu;
}

/* clip_global_norm, python/mxnet/gluon/utils.py:122-161:
 *   total = sqrt(sum_arrays dot(x,x)); scale = min(max_norm/(total+1e-8), 1); x *= scale  (all fp32 NDArray ops).
 * dot(x,x) on CPU NDArrays is a BLAS sdot in the reference (no BLAS here): we accumulate each array
 * in double and round to fp32, i.e. the correctly rounded value any accurate sdot approximates.
 * Returns the total norm. */
float oracle_clip_global_norm(int count, float** arrays, const int64_t* sizes, float max_norm,
                              int nthread) {
  float total_sq = 0.f;
  float* part = (float*)malloc(sizeof(float) * (size_t)count);
#pragma omp parallel for num_threads(nthread)
  for (int t = 0; t < count; ++t) {
    double s = 0;
    for (int64_t j = 0; j < sizes[t]; ++j) s += (double)arrays[t][j] * (double)arrays[t][j];
    part[t] = (float)s;
  }
  for (int t = 0; t < count; ++t) total_sq += part[t];   /* add_n of fp32 scalars */
  free(part);
  float total = sqrtf(total_sq);
  float scale = max_norm / (total + 1e-8f);
  if (scale > 1.0f) scale = 1.0f;
  for (int t = 0; t < count; ++t) {
#pragma omp parallel for num_threads(nthread)
    for (int64_t j = 0; j < sizes[t]; ++j) arrays[t][j] *= scale;
  }
  return total;
}
