/* CPU ORACLE (test / baseline infrastructure, NOT product code).
 *
 * Plain-C restatement of what Compute.scala executes on a CPU OpenCL device for BASELINE config C2,
 * `tanh(a * b + c).sum` with `c` broadcast per row (left-aligned, Tensors.scala:208-222):
 *   c2_map   = the generated `jit_kernel` for the tree Tanh(Plus(Times(a, b), c))  (OpenCLKernelBuilder.scala:135-221):
 *              one work-item per output element, scalar fp32 code, row-major output; the OpenCL CPU runtime spreads
 *              work-groups over all cores, restated here as a pthread split over rows.
 *   c2_sum16 = the sequential `reduce` program used when the device type is CPU (Tensors.scala:313-351, launched
 *              with global size 1, Tensors.scala:683-696): 16 strided lanes, hi/lo folds, then the tail.
 * Parity: pinned through oracle/tensors_oracle.py (same arithmetic; tests/test_oracle_c.py compares the two).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
 */
#include <math.h>
#include <pthread.h>
#include <stddef.h>
#include <unistd.h>

int c2_threads(void) {
  long n = sysconf(_SC_NPROCESSORS_ONLN);
  return n < 1 ? 1 : (n > 256 ? 256 : (int)n);
}

typedef struct {
  const float *a, *b, *c;
  float *y;
  size_t row0, row1, cols;
  int per_column;
} c2_job;

/* one rounding per node (no FMA contraction: compiled with -ffp-contract=off) */
static void *c2_worker(void *arg) {
  const c2_job *J = (const c2_job *)arg;
  for (size_t i = J->row0; i < J->row1; ++i) {
    const float *ai = J->a + i * J->cols, *bi = J->b + i * J->cols;
    float *yi = J->y + i * J->cols;
    for (size_t j = 0; j < J->cols; ++j) {
      float p = ai[j] * bi[j];
      float s = p + (J->per_column ? J->c[j] : J->c[i]);
      yi[j] = tanhf(s);
    }
  }
  return 0;
}

static void c2_map(const float *a, const float *b, const float *c, float *y, size_t rows, size_t cols, int per_column,
                   int threads) {
  if (threads < 1) threads = 1;
  if ((size_t)threads > rows) threads = rows ? (int)rows : 1;
  pthread_t tid[256];
  c2_job job[256];
  if (threads > 256) threads = 256;
  for (int t = 0; t < threads; ++t) {
    job[t].a = a; job[t].b = b; job[t].c = c; job[t].y = y; job[t].cols = cols; job[t].per_column = per_column;
    job[t].row0 = rows * (size_t)t / (size_t)threads;
    job[t].row1 = rows * (size_t)(t + 1) / (size_t)threads;
    if (t + 1 < threads) pthread_create(&tid[t], 0, c2_worker, &job[t]);
  }
  c2_worker(&job[threads - 1]);
  for (int t = 0; t + 1 < threads; ++t) pthread_join(tid[t], 0);
}

/* y[i][j] = tanh(a[i][j] * b[i][j] + c[i]) -- c broadcast per ROW (left-aligned auto-broadcast) */
void c2_map_rows(const float *a, const float *b, const float *c, float *y, size_t rows, size_t cols, int threads) {
  c2_map(a, b, c, y, rows, cols, 0, threads);
}

/* column variant: c.reshape([1, cols]) */
void c2_map_cols(const float *a, const float *b, const float *c, float *y, size_t rows, size_t cols, int threads) {
  c2_map(a, b, c, y, rows, cols, 1, threads);
}

/* Tensors.scala:313-351 with get_global_size(0) == 1 */
float c2_sum16(const float *x, size_t n) {
  float acc[16];
  const size_t vlen = n / 16;
  float total = 0.0f;
  if (vlen >= 1) {
    for (int l = 0; l < 16; ++l) acc[l] = x[l];
    for (size_t k = 1; k < vlen; ++k)
      for (int l = 0; l < 16; ++l) acc[l] = acc[l] + x[k * 16 + l];
    float f8[8], f4[4], f2[2];
    for (int l = 0; l < 8; ++l) f8[l] = acc[8 + l] + acc[l];
    for (int l = 0; l < 4; ++l) f4[l] = f8[4 + l] + f8[l];
    for (int l = 0; l < 2; ++l) f2[l] = f4[2 + l] + f4[l];
    total = f2[0] + f2[1];
  }
  for (size_t i = vlen * 16; i < n; ++i) total = total + x[i];
  return total;
}

/* ---- inputs: the reference's Wang-hash uniform generator (Tensors.scala:107-116 hash, :436-439 program):
 * buffer[i] = hash(i ^ seed) / 4294967296.0f, then `* mul + add` as two fp32 operations (bench.py maps to [-2, 2) with
 * mul = 4, add = -2 -- the same expression, hence the same bits, as the GPU arm's `random(shape, seed) * 4.0 - 2.0`). */
static unsigned c2_wang(unsigned v) {
  v = (v ^ 61u) ^ (v >> 16);
  v *= 9u;
  v ^= v << 4;
  v *= 0x27d4eb2du;
  v ^= v >> 15;
  return v;
}

typedef struct {
  float *out;
  size_t i0, i1;
  unsigned seed;
  float mul, add;
} c2_rng_job;

static void *c2_rng_worker(void *arg) {
  const c2_rng_job *J = (const c2_rng_job *)arg;
  for (size_t i = J->i0; i < J->i1; ++i) {
    float u = (float)c2_wang((unsigned)i ^ J->seed) / 4294967296.0f;
    float m = u * J->mul;
    J->out[i] = m + J->add;
  }
  return 0;
}

void c2_wang_uniform(float *out, size_t n, unsigned seed, float mul, float add, int threads) {
  if (threads < 1) threads = 1;
  if (threads > 256) threads = 256;
  pthread_t tid[256];
  c2_rng_job job[256];
  for (int t = 0; t < threads; ++t) {
    job[t].out = out; job[t].seed = seed; job[t].mul = mul; job[t].add = add;
    job[t].i0 = n * (size_t)t / (size_t)threads;
    job[t].i1 = n * (size_t)(t + 1) / (size_t)threads;
    if (t + 1 < threads) pthread_create(&tid[t], 0, c2_rng_worker, &job[t]);
  }
  c2_rng_worker(&job[threads - 1]);
  for (int t = 0; t + 1 < threads; ++t) pthread_join(tid[t], 0);
}

/* ---- best-effort CPU mode: what a tuned CPU implementation of the same expression would do --
 * map and sum fused (nothing materialised: 8 B/element instead of 16), every core, fp64 partial per thread.  NOT what
 * the reference executes (it materialises, then reduces on ONE work-item); reported beside the literal mode. */
typedef struct {
  const float *a, *b, *c;
  size_t row0, row1, cols;
  double partial;
} c2_fused_job;

static void *c2_fused_worker(void *arg) {
  c2_fused_job *J = (c2_fused_job *)arg;
  double acc = 0.0;
  for (size_t i = J->row0; i < J->row1; ++i) {
    const float *ai = J->a + i * J->cols, *bi = J->b + i * J->cols;
    const float ci = J->c[i];
    for (size_t j = 0; j < J->cols; ++j) {
      float p = ai[j] * bi[j];
      float s = p + ci;
      acc += (double)tanhf(s);
    }
  }
  J->partial = acc;
  return 0;
}

double c2_fused_sum(const float *a, const float *b, const float *c, size_t rows, size_t cols, int threads) {
  if (threads < 1) threads = 1;
  if (threads > 256) threads = 256;
  if ((size_t)threads > rows) threads = rows ? (int)rows : 1;
  pthread_t tid[256];
  c2_fused_job job[256];
  for (int t = 0; t < threads; ++t) {
    job[t].a = a; job[t].b = b; job[t].c = c; job[t].cols = cols; job[t].partial = 0.0;
    job[t].row0 = rows * (size_t)t / (size_t)threads;
    job[t].row1 = rows * (size_t)(t + 1) / (size_t)threads;
    if (t + 1 < threads) pthread_create(&tid[t], 0, c2_fused_worker, &job[t]);
  }
  c2_fused_worker(&job[threads - 1]);
  for (int t = 0; t + 1 < threads; ++t) pthread_join(tid[t], 0);
  double total = 0.0;
  for (int t = 0; t < threads; ++t) total += job[t].partial;
  return total;
}
