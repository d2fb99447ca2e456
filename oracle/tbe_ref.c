/*
 * TEST INFRASTRUCTURE ONLY — plain C restatement of the reference's CPU pooled lookup and
 * fused backward+optimizer, used (a) as the checker in tests/ and smoke(), (b) as the timed
 * `cpu_baseline` / `--impl reference` arm of bench.py.  The product never links this file.
 *
 * forward : third_party/fbgemm/fbgemm_gpu/codegen/embedding_forward_split_cpu.cpp:117-156
 *           (== EmbeddingSpMDM_ref, third_party/fbgemm/src/RefImplementations.cc:1186-1227)
 * backward: third_party/fbgemm/fbgemm_gpu/codegen/embedding_backward_split_cpu_template.cpp:193-224
 *           after a CSR->CSC transpose by (row, position) (batched_csr2csc,
 *           embedding_forward_split_cpu.cpp:322-564), with split_weight_update_cpu of
 *           rowwise_adagrad (embedding_backward_code_generator.py:402-415) and sgd (:633-637).
 * Parallelised with OpenMP the way the reference parallelises with at::parallel_for
 * (over the batch in forward, over tables / column blocks in backward).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* out[b, D_offsets[t] + d] = sum_p (w_p *) W_t[idx_p, d]; MEAN scales by 1/L for L > 0 */
void oracle_tbe_forward_f32(const float* weights, const int64_t* weights_offsets,
                            const int32_t* D_offsets, int32_t T, int32_t B, const int64_t* indices,
                            const int64_t* offsets, const float* psw, int mean, float* out) {
  const int64_t total_D = D_offsets[T];
#pragma omp parallel for schedule(static)
  for (int32_t b = 0; b < B; ++b) {
    for (int32_t t = 0; t < T; ++t) {
      const int32_t d0 = D_offsets[t], D = D_offsets[t + 1] - D_offsets[t];
      const float* tab = weights + weights_offsets[t];
      float* o = out + (int64_t)b * total_D + d0;
      const int64_t s = offsets[(int64_t)t * B + b], e = offsets[(int64_t)t * B + b + 1];
      for (int32_t d = 0; d < D; ++d) o[d] = 0.f;
      for (int64_t p = s; p < e; ++p) {
        const float* row = tab + indices[p] * D;
        const float w = psw ? psw[p] : 1.f;
        for (int32_t d = 0; d < D; ++d) o[d] = fmaf(w, row[d], o[d]);
      }
      if (mean && e > s) {
        const float sc = 1.0f / (float)(e - s);
        for (int32_t d = 0; d < D; ++d) o[d] *= sc;
      }
    }
  }
}

typedef struct {
  int64_t key; /* linear (table,row) key */
  int32_t t;   /* feature */
  int32_t b;
  int64_t p; /* original position: ties keep the stable t-major, b, in-bag order */
} occ_t;

static int occ_cmp(const void* a, const void* b) {
  const occ_t* x = (const occ_t*)a;
  const occ_t* y = (const occ_t*)b;
  if (x->key != y->key) return x->key < y->key ? -1 : 1;
  return x->p < y->p ? -1 : (x->p > y->p ? 1 : 0);
}

/* optimizer: 0 = EXACT_SGD, 1 = EXACT_ROWWISE_ADAGRAD.  In-place on weights / momentum1. */
void oracle_tbe_backward_f32(const float* grad, int64_t grad_stride, float* weights,
                             const int64_t* weights_offsets, const int32_t* D_offsets,
                             const int64_t* hash_size_cumsum, int32_t T, int32_t B,
                             const int64_t* indices, const int64_t* offsets, const float* psw,
                             int mean, int optimizer, float* momentum1,
                             const int64_t* momentum1_offsets, float lr, float eps) {
  const int64_t N = offsets[(int64_t)T * B];
  if (N == 0) return;
  occ_t* occ = (occ_t*)malloc(sizeof(occ_t) * (size_t)N);
#pragma omp parallel for schedule(static)
  for (int64_t bt = 0; bt < (int64_t)T * B; ++bt) {
    const int32_t t = (int32_t)(bt / B), b = (int32_t)(bt % B);
    for (int64_t p = offsets[bt]; p < offsets[bt + 1]; ++p) {
      occ[p].key = hash_size_cumsum[t] + indices[p];
      occ[p].t = t;
      occ[p].b = b;
      occ[p].p = p;
    }
  }
  qsort(occ, (size_t)N, sizeof(occ_t), occ_cmp);
  /* run starts */
  int64_t* starts = (int64_t*)malloc(sizeof(int64_t) * (size_t)(N + 1));
  int64_t U = 0;
  for (int64_t i = 0; i < N; ++i)
    if (i == 0 || occ[i].key != occ[i - 1].key) starts[U++] = i;
  starts[U] = N;
#pragma omp parallel
  {
    float* g = (float*)malloc(sizeof(float) * 4096);
#pragma omp for schedule(dynamic, 64)
    for (int64_t u = 0; u < U; ++u) {
      const occ_t* first = &occ[starts[u]];
      const int32_t t0 = first->t;
      const int32_t D = D_offsets[t0 + 1] - D_offsets[t0];
      memset(g, 0, sizeof(float) * (size_t)D);
      for (int64_t i = starts[u]; i < starts[u + 1]; ++i) {
        const occ_t* o = &occ[i];
        const float* row = grad + (int64_t)o->b * grad_stride + D_offsets[o->t];
        float sc = 1.f;
        if (mean) {
          const int64_t bt = (int64_t)o->t * B + o->b;
          sc = 1.0f / (float)(offsets[bt + 1] - offsets[bt]);
        }
        if (psw) sc *= psw[o->p];
        for (int32_t d = 0; d < D; ++d) g[d] += sc * row[d];
      }
      const int64_t idx = first->key - hash_size_cumsum[t0];
      float* w = weights + weights_offsets[t0] + idx * D;
      if (optimizer == 1) {
        float ss = 0.f;
        for (int32_t d = 0; d < D; ++d) ss += g[d] * g[d];
        float* m = momentum1 + momentum1_offsets[t0] + idx;
        const float mnew = *m + ss / (float)D;
        *m = mnew;
        const float mult = lr / (sqrtf(mnew) + eps);
        for (int32_t d = 0; d < D; ++d) w[d] -= g[d] * mult;
      } else {
        for (int32_t d = 0; d < D; ++d) w[d] -= lr * g[d];
      }
    }
    free(g);
  }
  free(starts);
  free(occ);
}

/* InteractionArch.forward (torchrec/models/dlrm.py:193-219), one sample per iteration */
void oracle_interaction_forward(const float* dense, const float* sparse, int32_t B, int32_t F,
                                int32_t D, float* out) {
  const int n = F + 1;
  const int64_t out_dim = D + (int64_t)n * (n - 1) / 2;
#pragma omp parallel for schedule(static)
  for (int32_t s = 0; s < B; ++s) {
    float* o = out + s * out_dim;
    const float* x0 = dense + (int64_t)s * D;
    const float* xs = sparse + (int64_t)s * F * D;
    memcpy(o, x0, sizeof(float) * (size_t)D);
    int64_t k = D;
    for (int i = 0; i < n; ++i) {
      const float* xi = i == 0 ? x0 : xs + (int64_t)(i - 1) * D;
      for (int j = i + 1; j < n; ++j) {
        const float* xj = xs + (int64_t)(j - 1) * D;
        float acc = 0.f;
        for (int d = 0; d < D; ++d) acc += xi[d] * xj[d];
        o[k++] = acc;
      }
    }
  }
}
