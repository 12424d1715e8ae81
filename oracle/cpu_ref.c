/*
 * cpu_ref.c -- C restatement of the reference's per-block algorithm, for the timed CPU
 * baseline (bench.py `cpu_baseline` and `--impl reference`) and as a second checker.
 * TEST INFRASTRUCTURE ONLY: the product never links or loads this file.
 *
 * The reference (yupbank/tensorframes) has no native code of its own; a block goes through
 * these steps in the JVM + TensorFlow-CPU (paths under
 * /root/reference/src/main/scala/org/tensorframes/impl/):
 *
 *   (1) TFDataOps.convert -> DataOps.convertFast0 -> TensorConverter.append
 *       (TFDataOps.scala:27-59, DataOps.scala:63-81, datatypes.scala:114-127, 340-342):
 *       every cell scalar is appended one by one into a freshly reserved heap buffer.
 *   (2) Tensor.create(shape, buffer) (datatypes.scala:344-347): the heap buffer is copied
 *       into a native tensor.
 *   (3) Session.run (DebugRowOps.scala:794 / 912): one TF-CPU kernel per graph node, each
 *       writing a full output tensor (no fusion across nodes).
 *   (4) Tensor.writeTo(buffer) after a NaN/0 prefill (datatypes.scala:365-370): the result is
 *       copied back to the JVM heap.
 *   (5) DataOps.convertBackFast0 / getColumnFast0 (DataOps.scala:20-61, 146-166): one output
 *       row object per row.
 *   Partitions run as independent Spark tasks (local[N] = N task threads,
 *   DebugRowOps.scala:377-391); reduce_blocks then merges the per-partition rows pairwise
 *   by running the graph again on 2-row blocks (DebugRowOps.scala:741-750, 524-525).
 *
 * This file restates exactly those steps on data that is ALREADY columnar and unboxed, one
 * OpenMP thread per partition.  It therefore leaves out the JVM costs the reference also
 * pays (boxed Row access, GenericRow allocation, Graph+Session construction per block) and
 * is an upper bound on the reference's speed.
 *
 * Build: gcc -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp -fPIC -shared (see
 * __graft_entry__.build_oracle).  -ffp-contract=off keeps x*a+b as two roundings, like the
 * separate Mul and Add kernels of TF.
 */
#include <math.h>
#include <omp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static void part_range(int64_t n, int parts, int p, int64_t* lo, int64_t* hi) {
  *lo = n * p / parts;
  *hi = n * (p + 1) / parts;
}

/* steps (1)+(2): append cell scalars one at a time, then copy the buffer into a "tensor" */
static double* pack_block_f64(const double* col, int64_t rows, int cell) {
  const int64_t n = rows * cell;
  double* heap = (double*)malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
  int64_t w = 0;
  for (int64_t r = 0; r < rows; ++r)
    for (int c = 0; c < cell; ++c) heap[w++] = col[r * cell + c];
  double* tensor = (double*)malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
  memcpy(tensor, heap, (size_t)n * sizeof(double));
  free(heap);
  return tensor;
}

static float* pack_block_f32(const float* col, int64_t rows, int cell) {
  const int64_t n = rows * cell;
  float* heap = (float*)malloc((size_t)(n > 0 ? n : 1) * sizeof(float));
  int64_t w = 0;
  for (int64_t r = 0; r < rows; ++r)
    for (int c = 0; c < cell; ++c) heap[w++] = col[r * cell + c];
  float* tensor = (float*)malloc((size_t)(n > 0 ? n : 1) * sizeof(float));
  memcpy(tensor, heap, (size_t)n * sizeof(float));
  free(heap);
  return tensor;
}

/* TF Sum / Min along axis 0 of a [rows, cell] tensor */
static void sum_axis0_f64(const double* t, int64_t rows, int cell, double* out) {
  for (int c = 0; c < cell; ++c) out[c] = 0.0;
  for (int64_t r = 0; r < rows; ++r)
    for (int c = 0; c < cell; ++c) out[c] += t[r * cell + c];
}
static void min_axis0_f64(const double* t, int64_t rows, int cell, double* out) {
  for (int c = 0; c < cell; ++c) out[c] = INFINITY;
  for (int64_t r = 0; r < rows; ++r)
    for (int c = 0; c < cell; ++c) {
      double v = t[r * cell + c];
      out[c] = (v < out[c]) ? v : out[c];
    }
}

/*
 * BASELINE config 2: reduce_blocks with y = Sum(y_input, [0]), z = Min(z_input, [0]) over two
 * f64[rows, cell] columns.  performReduceBlock per partition (DebugRowOps.scala:872-895),
 * then the pairwise merge.  out_sum / out_min: [cell].
 */
void ref_reduce_blocks_sum_min_f64(const double* y, const double* z, int64_t rows, int cell, int partitions,
                                   int threads, double* out_sum, double* out_min) {
  double* psum = (double*)malloc((size_t)partitions * cell * sizeof(double));
  double* pmin = (double*)malloc((size_t)partitions * cell * sizeof(double));
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
  for (int p = 0; p < partitions; ++p) {
    int64_t lo, hi;
    part_range(rows, partitions, p, &lo, &hi);
    double* ty = pack_block_f64(y + lo * cell, hi - lo, cell);   /* (1)+(2) */
    double* tz = pack_block_f64(z + lo * cell, hi - lo, cell);
    double* ry = (double*)malloc((size_t)cell * sizeof(double));  /* (3) output tensors */
    double* rz = (double*)malloc((size_t)cell * sizeof(double));
    sum_axis0_f64(ty, hi - lo, cell, ry);
    min_axis0_f64(tz, hi - lo, cell, rz);
    memcpy(psum + (size_t)p * cell, ry, (size_t)cell * sizeof(double));  /* (4)+(5): one row */
    memcpy(pmin + (size_t)p * cell, rz, (size_t)cell * sizeof(double));
    free(ty); free(tz); free(ry); free(rz);
  }
  /* RDD.reduce(reducePairBlock): fold partition rows two at a time through the same graph */
  double* acc_s = (double*)malloc((size_t)2 * cell * sizeof(double));
  double* acc_m = (double*)malloc((size_t)2 * cell * sizeof(double));
  memcpy(acc_s, psum, (size_t)cell * sizeof(double));
  memcpy(acc_m, pmin, (size_t)cell * sizeof(double));
  for (int p = 1; p < partitions; ++p) {
    memcpy(acc_s + cell, psum + (size_t)p * cell, (size_t)cell * sizeof(double));
    memcpy(acc_m + cell, pmin + (size_t)p * cell, (size_t)cell * sizeof(double));
    double* ts = pack_block_f64(acc_s, 2, cell);
    double* tm = pack_block_f64(acc_m, 2, cell);
    sum_axis0_f64(ts, 2, cell, acc_s);
    min_axis0_f64(tm, 2, cell, acc_m);
    free(ts); free(tm);
  }
  memcpy(out_sum, acc_s, (size_t)cell * sizeof(double));
  memcpy(out_min, acc_m, (size_t)cell * sizeof(double));
  free(acc_s); free(acc_m); free(psum); free(pmin);
}

/*
 * BASELINE config 3: map_blocks with z = Cast_f32(Maximum(Add(Mul(x, a), b), 0)) on an f64
 * column.  performMap per partition (DebugRowOps.scala:766-803): pack, four TF kernels with
 * three materialised intermediates, writeTo, one output row per input row.
 */
void ref_map_blocks_chain_f64(const double* x, int64_t rows, int partitions, int threads, double a, double b,
                              float* out) {
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
  for (int p = 0; p < partitions; ++p) {
    int64_t lo, hi;
    part_range(rows, partitions, p, &lo, &hi);
    const int64_t n = hi - lo;
    double* tx = pack_block_f64(x + lo, n, 1);                       /* (1)+(2) */
    double* t1 = (double*)malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
    double* t2 = (double*)malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
    double* t3 = (double*)malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
    float* t4 = (float*)malloc((size_t)(n > 0 ? n : 1) * sizeof(float));
    for (int64_t i = 0; i < n; ++i) t1[i] = tx[i] * a;               /* (3) Mul      */
    for (int64_t i = 0; i < n; ++i) t2[i] = t1[i] + b;               /*     Add      */
    for (int64_t i = 0; i < n; ++i) t3[i] = (t2[i] < 0.0) ? 0.0 : t2[i];  /* Maximum */
    for (int64_t i = 0; i < n; ++i) t4[i] = (float)t3[i];            /*     Cast     */
    float* heap = (float*)malloc((size_t)(n > 0 ? n : 1) * sizeof(float));
    for (int64_t i = 0; i < n; ++i) heap[i] = NAN;                   /* (4) prefill + writeTo */
    memcpy(heap, t4, (size_t)n * sizeof(float));
    for (int64_t i = 0; i < n; ++i) out[lo + i] = heap[i];           /* (5) one row per row */
    free(tx); free(t1); free(t2); free(t3); free(t4); free(heap);
  }
}

/*
 * BASELINE config 4: groupBy(int64 key).aggregate(Sum(v_input, [0])) over f32[rows, cell].
 * The UDAF buffers rows per key and block-reduces them every bufferSize rows
 * (DebugRowOps.scala:601-695); numerically that is a running per-key sum in arrival order,
 * compacted in groups of at most 11 rows.  Restated here as an open-addressing hash table
 * per partition with a per-key pending count, then a merge of the partition tables.
 * out_keys: [max_groups], out: [max_groups, cell]; returns the number of groups (or -1 if
 * max_groups is too small).
 */
typedef struct { int64_t key; int used; int pending; float* acc; float* buf; } slot_t;

static uint64_t mix64(uint64_t x) {
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL; x ^= x >> 27; x *= 0x94d049bb133111ebULL; x ^= x >> 31;
  return x;
}

int64_t ref_aggregate_sum_f32(const int64_t* keys, const float* v, int64_t rows, int cell, int partitions,
                              int threads, int64_t max_groups, int64_t* out_keys, float* out) {
  int64_t cap = 16;
  while (cap < 2 * max_groups) cap <<= 1;
  slot_t** tables = (slot_t**)calloc((size_t)partitions, sizeof(slot_t*));
  const int B = 10;  /* TensorFlowUDAF bufferSize (DebugRowOps.scala:573) */
  int overflow = 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
  for (int p = 0; p < partitions; ++p) {
    int64_t lo, hi;
    part_range(rows, partitions, p, &lo, &hi);
    slot_t* tab = (slot_t*)calloc((size_t)cap, sizeof(slot_t));
    tables[p] = tab;
    int64_t used = 0;
    for (int64_t r = lo; r < hi; ++r) {
      int64_t k = keys[r];
      uint64_t h = mix64((uint64_t)k) & (uint64_t)(cap - 1);
      while (tab[h].used && tab[h].key != k) h = (h + 1) & (uint64_t)(cap - 1);
      slot_t* s = &tab[h];
      if (!s->used) {
        if (++used > max_groups) { overflow = 1; break; }
        s->used = 1; s->key = k; s->pending = 0;
        s->acc = (float*)calloc((size_t)cell, sizeof(float));
        s->buf = (float*)malloc((size_t)(B + 1) * cell * sizeof(float));
      }
      memcpy(s->buf + (size_t)s->pending * cell, v + r * cell, (size_t)cell * sizeof(float));  /* input.copy() */
      if (++s->pending > B) {   /* compact: block reduce of the buffered rows */
        for (int c = 0; c < cell; ++c) {
          float t = 0.f;
          for (int j = 0; j < s->pending; ++j) t += s->buf[(size_t)j * cell + c];
          s->buf[c] = t;
        }
        s->pending = 1;
      }
    }
    for (int64_t h = 0; h < cap; ++h) {
      slot_t* s = &tab[h];
      if (!s->used) continue;
      for (int c = 0; c < cell; ++c) {
        float t = 0.f;
        for (int j = 0; j < s->pending; ++j) t += s->buf[(size_t)j * cell + c];
        s->acc[c] = t;
      }
    }
  }
  /* merge partition tables (Spark's shuffle + final merge) */
  int64_t groups = 0;
  if (!overflow) {
    slot_t* fin = (slot_t*)calloc((size_t)cap, sizeof(slot_t));
    for (int p = 0; p < partitions && !overflow; ++p) {
      slot_t* tab = tables[p];
      for (int64_t i = 0; i < cap; ++i) {
        if (!tab[i].used) continue;
        uint64_t h = mix64((uint64_t)tab[i].key) & (uint64_t)(cap - 1);
        while (fin[h].used && fin[h].key != tab[i].key) h = (h + 1) & (uint64_t)(cap - 1);
        if (!fin[h].used) {
          if (groups >= max_groups) { overflow = 1; break; }
          fin[h].used = 1; fin[h].key = tab[i].key; fin[h].pending = (int)groups;
          out_keys[groups] = tab[i].key;
          memcpy(out + (size_t)groups * cell, tab[i].acc, (size_t)cell * sizeof(float));
          ++groups;
        } else {
          float* dst = out + (size_t)fin[h].pending * cell;
          for (int c = 0; c < cell; ++c) dst[c] += tab[i].acc[c];
        }
      }
    }
    free(fin);
  }
  for (int p = 0; p < partitions; ++p) {
    if (!tables[p]) continue;
    for (int64_t i = 0; i < cap; ++i)
      if (tables[p][i].used) { free(tables[p][i].acc); free(tables[p][i].buf); }
    free(tables[p]);
  }
  free(tables);
  return overflow ? -1 : groups;
}

/*
 * BASELINE config 5: map_blocks with MatMul(x[rows,k], w[k,n]) + bias[n], float32.
 * TF-CPU's MatMul is an Eigen contraction in float32; restated as a straightforward
 * i-k-j float32 product followed by the separate Add kernel.
 */
void ref_map_blocks_matmul_bias_f32(const float* x, const float* w, const float* bias, int64_t rows, int k, int n,
                                    int partitions, int threads, float* out) {
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
  for (int p = 0; p < partitions; ++p) {
    int64_t lo, hi;
    part_range(rows, partitions, p, &lo, &hi);
    const int64_t m = hi - lo;
    float* tx = pack_block_f32(x + lo * k, m, k);
    float* t1 = (float*)calloc((size_t)(m > 0 ? m : 1) * n, sizeof(float));
    for (int64_t i = 0; i < m; ++i)
      for (int kk = 0; kk < k; ++kk) {
        const float xv = tx[i * k + kk];
        const float* wr = w + (size_t)kk * n;
        float* o = t1 + i * n;
        for (int j = 0; j < n; ++j) o[j] += xv * wr[j];
      }
    float* t2 = (float*)malloc((size_t)(m > 0 ? m : 1) * n * sizeof(float));
    for (int64_t i = 0; i < m; ++i)
      for (int j = 0; j < n; ++j) t2[i * n + j] = t1[i * n + j] + bias[j];
    memcpy(out + lo * n, t2, (size_t)m * n * sizeof(float));
    free(tx); free(t1); free(t2);
  }
}

int ref_max_threads(void) { return omp_get_max_threads(); }
