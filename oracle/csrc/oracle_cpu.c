/*
 * oracle_cpu.c -- plain-C restatement of the decode hot path.  TEST / BASELINE
 * INFRASTRUCTURE ONLY: used by tests/ as a second checker and by bench.py as the
 * timed CPU baseline ("kind": "port").  Never linked into the product library.
 *
 * parity: attention/KV-write/metadata arithmetic is pinned against the reference's
 * fixtures through the NumPy oracle (tests/test_oracle_golden.py compares this file
 * with oracle/*.py); the affine-quantized arithmetic is "parity unpinned" (MLX 0.32.0
 * is a third-party dependency absent from /root/reference, see oracle/affine_quant.py).
 *
 * Reference files each function follows:
 *   oracle_reshape_and_cache  vllm_metal/metal/kernels_v2/reshape_and_cache.metal:60-95
 *   oracle_paged_attention    vllm_metal/metal/kernels_v2/pagedattention.metal:913-926 (eff ctx),
 *                             :1129-1138 (mask, log2 domain), :1318-1328 (sink), :1383 (1/(l+1e-6));
 *                             tools/attention_bench_utils.py:11-62 (gather by block table)
 *   oracle_qmatmul            vllm_metal/gguf/mlx_native.py:212-222 (mx.quantized_matmul,
 *                             transpose=True) -- MLX affine: w' = s*q + b, groups along K
 *   oracle_dequantize         vllm_metal/gguf/mlx_native.py:232-239 (mx.dequantize)
 *   oracle_rms_norm/rope/silu_mul/argmax: MLX built-ins around sdpa_forward
 *                             (attention/impls/sdpa.py:348-368, v1/sampling_batch.py:342-348)
 *
 * 16-bit tensors are bf16 bit patterns (uint16_t); math is fp32; one RNE rounding
 * where the path stores a tensor.  Build: gcc -O3 -fopenmp -shared -fPIC (oracle/build.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

static inline float bf2f(uint16_t v) {
  uint32_t u = (uint32_t)v << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  /* NaN */
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* Pin the OpenMP team size (bench.py: all host cores, whatever OMP_NUM_THREADS a launcher injected). */
void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* cache[slot / bs][slot % bs] = row t; negative slots skipped. row_elems = H_kv * D. */
void oracle_reshape_and_cache(const uint16_t* key, const uint16_t* value, uint16_t* key_cache,
                              uint16_t* value_cache, const int64_t* slot_mapping, int num_tokens,
                              int row_elems) {
  for (int t = 0; t < num_tokens; ++t) {
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue;
    memcpy(key_cache + slot * row_elems, key + (int64_t)t * row_elems, (size_t)row_elems * 2);
    memcpy(value_cache + slot * row_elems, value + (int64_t)t * row_elems, (size_t)row_elems * 2);
  }
}

/* Paged varlen attention, bf16 in/out, fp32 math.  Layouts as include/b200_paged_ops.h:
 * q/out [T, Hq, D]; caches [num_blocks, block_size, Hkv, D]; block_tables [S, max_blocks];
 * seq_lens [S]; cu_seqlens_q [S+1]; sinks [Hq] or NULL; sliding_window < 0 disables.
 * Rows with no visible KV are left untouched. */
void oracle_paged_attention(uint16_t* out, const uint16_t* q, const uint16_t* key_cache,
                            const uint16_t* value_cache, const int32_t* block_tables,
                            const int32_t* seq_lens, const int32_t* cu_seqlens_q, const float* sinks,
                            int num_seqs, int num_q_heads, int num_kv_heads, int head_dim,
                            int block_size, int max_blocks, int sliding_window, float scale,
                            float softcap) {
  const int G = num_q_heads / num_kv_heads;
  const int D = head_dim;
  const float LOG2E = 1.4426950408889634f;
  const int64_t tok_stride = (int64_t)num_kv_heads * D;
  int max_len = 1;
  for (int s = 0; s < num_seqs; ++s)
    if (seq_lens[s] > max_len) max_len = seq_lens[s];
  /* one task per (row, kv head) */
  const int T = cu_seqlens_q[num_seqs];
#pragma omp parallel
  {
    float* sc = (float*)malloc(sizeof(float) * (size_t)max_len * G);
    float* acc = (float*)malloc(sizeof(float) * (size_t)G * D);
    float* qf = (float*)malloc(sizeof(float) * (size_t)G * D);
#pragma omp for schedule(dynamic, 1) collapse(2)
    for (int t = 0; t < T; ++t) {
      for (int kvh = 0; kvh < num_kv_heads; ++kvh) {
        int s = 0;
        while (cu_seqlens_q[s + 1] <= t) ++s;
        const int q_len = cu_seqlens_q[s + 1] - cu_seqlens_q[s];
        const int eff = seq_lens[s] - q_len + (t - cu_seqlens_q[s]) + 1;
        if (eff <= 0) continue;
        int lo = 0;
        if (sliding_window >= 0) lo = eff - sliding_window > 0 ? eff - sliding_window : 0;
        const int n = eff - lo;
        const int32_t* table = block_tables + (int64_t)s * max_blocks;
        for (int g = 0; g < G; ++g)
          for (int d = 0; d < D; ++d)
            qf[g * D + d] = bf2f(q[((int64_t)t * num_q_heads + kvh * G + g) * D + d]);
        /* scores */
        for (int j = 0; j < n; ++j) {
          const int pos = lo + j;
          const uint16_t* kr = key_cache +
                               ((int64_t)table[pos / block_size] * block_size + pos % block_size) * tok_stride +
                               (int64_t)kvh * D;
          float kf[512];
          for (int d = 0; d < D; ++d) kf[d] = bf2f(kr[d]);
          for (int g = 0; g < G; ++g) {
            float dot = 0.f;
            for (int d = 0; d < D; ++d) dot += qf[g * D + d] * kf[d];
            dot *= scale;
            if (softcap > 0.f) dot = tanhf(dot / softcap) * softcap;
            sc[(size_t)g * max_len + j] = dot * LOG2E;
          }
        }
        for (int g = 0; g < G; ++g) {
          float* sg = sc + (size_t)g * max_len;
          const int h = kvh * G + g;
          float m = -INFINITY;
          for (int j = 0; j < n; ++j) m = sg[j] > m ? sg[j] : m;
          float sk = 0.f;
          if (sinks) {
            sk = sinks[h] * LOG2E;
            m = sk > m ? sk : m;
          }
          if (!(m > -INFINITY)) m = 0.f;
          float l = 0.f;
          for (int j = 0; j < n; ++j) {
            sg[j] = exp2f(sg[j] - m);
            l += sg[j];
          }
          if (sinks) l += exp2f(sk - m);
          float* ag = acc + g * D;
          for (int d = 0; d < D; ++d) ag[d] = 0.f;
          for (int j = 0; j < n; ++j) {
            const int pos = lo + j;
            const uint16_t* vr = value_cache +
                                 ((int64_t)table[pos / block_size] * block_size + pos % block_size) * tok_stride +
                                 (int64_t)kvh * D;
            const float w = sg[j];
            for (int d = 0; d < D; ++d) ag[d] += w * bf2f(vr[d]);
          }
          const float inv = 1.f / (l + 1e-6f);
          uint16_t* orow = out + ((int64_t)t * num_q_heads + h) * D;
          for (int d = 0; d < D; ++d) orow[d] = f2bf(ag[d] * inv);
        }
      }
    }
    free(sc);
    free(acc);
    free(qf);
  }
}

static inline void dequant_row(float* dst, const uint32_t* w, const uint16_t* scales,
                               const uint16_t* biases, int K, int group, int bits) {
  const int per = 32 / bits;
  const uint32_t mask = (1u << bits) - 1u;
  for (int k0 = 0; k0 < K; k0 += per) {
    const uint32_t wd = w[k0 / per];
    const float s = bf2f(scales[k0 / group]), b = bf2f(biases[k0 / group]);
    for (int e = 0; e < per; ++e) dst[k0 + e] = s * (float)((wd >> (e * bits)) & mask) + b;
  }
}

/* out[r, :] = bf16(s*q+b) of row ids[r] (ids NULL -> r). */
void oracle_dequantize(uint16_t* out, const uint32_t* w, const uint16_t* scales, const uint16_t* biases,
                       const int32_t* ids, int num_rows, int K, int group, int bits) {
  const int words = K * bits / 32, groups = K / group;
#pragma omp parallel
  {
    float* tmp = (float*)malloc(sizeof(float) * (size_t)K);
#pragma omp for
    for (int r = 0; r < num_rows; ++r) {
      const int64_t row = ids ? ids[r] : r;
      dequant_row(tmp, w + row * words, scales + row * groups, biases + row * groups, K, group, bits);
      for (int k = 0; k < K; ++k) out[(int64_t)r * K + k] = f2bf(tmp[k]);
    }
    free(tmp);
  }
}

/* y[m,n] = bf16( sum_k x[m,k] * (s*q+b)[n,k] ), x bf16 [M,K], y bf16 [M,N]. */
void oracle_qmatmul(uint16_t* y, const uint16_t* x, const uint32_t* w, const uint16_t* scales,
                    const uint16_t* biases, int M, int N, int K, int group, int bits) {
  const int words = K * bits / 32, groups = K / group;
  float* xf = (float*)malloc(sizeof(float) * (size_t)M * K);
  for (int64_t i = 0; i < (int64_t)M * K; ++i) xf[i] = bf2f(x[i]);
#pragma omp parallel
  {
    float* wr = (float*)malloc(sizeof(float) * (size_t)K);
#pragma omp for schedule(static)
    for (int n = 0; n < N; ++n) {
      dequant_row(wr, w + (int64_t)n * words, scales + (int64_t)n * groups, biases + (int64_t)n * groups, K,
                  group, bits);
      for (int m = 0; m < M; ++m) {
        const float* xr = xf + (int64_t)m * K;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        for (int k = 0; k < K; k += 4) {
          a0 += xr[k] * wr[k];
          a1 += xr[k + 1] * wr[k + 1];
          a2 += xr[k + 2] * wr[k + 2];
          a3 += xr[k + 3] * wr[k + 3];
        }
        y[(int64_t)m * N + n] = f2bf((a0 + a1) + (a2 + a3));
      }
    }
    free(wr);
  }
  free(xf);
}

/* out = bf16(x * rsqrt(mean(x^2)+eps) * w); if residual: residual = bf16(residual + x) first. */
void oracle_rms_norm(uint16_t* out, const uint16_t* x, uint16_t* residual, const uint16_t* weight,
                     int rows, int dim, float eps) {
#pragma omp parallel for
  for (int r = 0; r < rows; ++r) {
    float ss = 0.f;
    for (int d = 0; d < dim; ++d) {
      float v = bf2f(x[(int64_t)r * dim + d]);
      if (residual) {
        v = bf2f(f2bf(v + bf2f(residual[(int64_t)r * dim + d])));
        residual[(int64_t)r * dim + d] = f2bf(v);
      }
      ss += v * v;
    }
    const float inv = 1.f / sqrtf(ss / (float)dim + eps);
    for (int d = 0; d < dim; ++d) {
      const float v = bf2f(residual ? residual[(int64_t)r * dim + d] : x[(int64_t)r * dim + d]);
      out[(int64_t)r * dim + d] = f2bf(v * inv * (weight ? bf2f(weight[d]) : 1.f));
    }
  }
}

/* in-place half-split RoPE on [T, H, D] at positions[T]. */
void oracle_rope(uint16_t* x, const int32_t* positions, int T, int H, int D, float base) {
  const int half = D / 2;
#pragma omp parallel for
  for (int t = 0; t < T; ++t) {
    for (int i = 0; i < half; ++i) {
      const double inv_freq = pow((double)base, -2.0 * i / D);
      const float cs = (float)cos(positions[t] * inv_freq), sn = (float)sin(positions[t] * inv_freq);
      for (int h = 0; h < H; ++h) {
        uint16_t* p = x + ((int64_t)t * H + h) * D;
        const float a = bf2f(p[i]), b = bf2f(p[i + half]);
        p[i] = f2bf(a * cs - b * sn);
        p[i + half] = f2bf(b * cs + a * sn);
      }
    }
  }
}

void oracle_silu_mul(uint16_t* out, const uint16_t* gate_up, int rows, int inter) {
#pragma omp parallel for
  for (int r = 0; r < rows; ++r)
    for (int i = 0; i < inter; ++i) {
      const float g = bf2f(gate_up[(int64_t)r * 2 * inter + i]);
      const float u = bf2f(gate_up[(int64_t)r * 2 * inter + inter + i]);
      const float s = bf2f(f2bf(g / (1.f + expf(-g))));
      out[(int64_t)r * inter + i] = f2bf(s * u);
    }
}

void oracle_argmax(int32_t* out, const uint16_t* logits, int rows, int vocab) {
#pragma omp parallel for
  for (int r = 0; r < rows; ++r) {
    float best = -INFINITY;
    int bi = 0;
    for (int i = 0; i < vocab; ++i) {
      const float v = bf2f(logits[(int64_t)r * vocab + i]);
      if (v > best) {
        best = v;
        bi = i;
      }
    }
    out[r] = bi;
  }
}
