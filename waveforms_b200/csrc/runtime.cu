// runtime.cu -- error text, device check, LFSR host math (mask table, GF(2) jump).
#include "common.cuh"
#include "lfsr_host.h"
#include <string.h>
#include <mutex>

namespace wf {

static thread_local char g_err[512] = "";

char* err_buf() { return g_err; }

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int cuda_fail(cudaError_t e, const char* where) {
  snprintf(g_err, sizeof(g_err), "CUDA error at %s: %s", where, cudaGetErrorString(e));
  return WF_ERR_CUDA;
}

static int g_dev_ok[64];   // 0 unknown, 1 ok, -1 bad
static int g_sm_count[64];

int require_device() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    cudaGetLastError();
    return fail(WF_ERR_DEVICE, "no CUDA device available (%s); libwfb200 has no CPU fallback",
                e == cudaSuccess ? "bad ordinal" : cudaGetErrorString(e));
  }
  if (g_dev_ok[dev] == 0) {
    cudaDeviceProp p;
    e = cudaGetDeviceProperties(&p, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceProperties");
    g_sm_count[dev] = p.multiProcessorCount;
    g_dev_ok[dev] = (p.major == 10) ? 1 : -1;
  }
  if (g_dev_ok[dev] < 0)
    return fail(WF_ERR_DEVICE, "device %d is not compute capability 10.x; libwfb200 is built for sm_100a only", dev);
  return WF_OK;
}

int sm_count() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  return g_sm_count[dev] > 0 ? g_sm_count[dev] : 148;
}

// ---------------------------------------------------------------------------
// XAPP052 maximal-length tap sets, one string per degree 2..64 (same
// polynomials as waveforms/glfsr/pn.py:6-72, re-encoded).
// ---------------------------------------------------------------------------
static const char* const kTaps[65] = {
    nullptr, nullptr,
    "2,1", "3,2", "4,3", "5,3", "6,5", "7,6", "8,6,5,4", "9,5", "10,7", "11,9",
    "12,6,4,1", "13,4,3,1", "14,5,3,1", "15,14", "16,15,13,4", "17,14", "18,11",
    "19,6,2,1", "20,17", "21,19", "22,21", "23,18", "24,23,22,17", "25,22",
    "26,6,2,1", "27,5,2,1", "28,25", "29,27", "30,6,4,1", "31,28", "32,22,2,1",
    "33,20", "34,27,2,1", "35,33", "36,25", "37,5,4,3,2,1", "38,6,5,1", "39,35",
    "40,38,21,19", "41,38", "42,41,20,19", "43,42,38,37", "44,43,18,17",
    "45,44,42,41", "46,45,26,25", "47,42", "48,47,21,20", "49,40", "50,49,24,23",
    "51,50,36,35", "52,49", "53,52,38,37", "54,53,18,17", "55,31", "56,55,35,34",
    "57,50", "58,39", "59,58,38,37", "60,59", "61,60,46,45", "62,61,6,5", "63,62",
    "64,63,61,60"};

uint64_t lfsr_mask(int degree) {
  if (degree < 2 || degree > 64) return 0;
  uint64_t m = 0;
  const char* s = kTaps[degree];
  while (*s) {
    int t = 0;
    while (*s >= '0' && *s <= '9') t = t * 10 + (*s++ - '0');
    m |= 1ull << (t - 1);
    if (*s == ',') ++s;
  }
  return m;
}

// One Galois step (glfsr.py:15-19) is linear over GF(2): column i of the step
// matrix is the image of unit vector e_i.
void gf2_step_matrix(uint64_t mask, int degree, Gf2Mat& a) {
  memset(&a, 0, sizeof(a));
  a.col[0] = mask;
  for (int i = 1; i < degree; ++i) a.col[i] = 1ull << (i - 1);
}

uint64_t gf2_apply(const Gf2Mat& a, int degree, uint64_t v) {
  uint64_t r = 0;
  for (int i = 0; i < degree; ++i)
    if ((v >> i) & 1) r ^= a.col[i];
  return r;
}

void gf2_mul(const Gf2Mat& a, const Gf2Mat& b, int degree, Gf2Mat& c) {
  Gf2Mat t;
  memset(&t, 0, sizeof(t));
  for (int i = 0; i < degree; ++i) t.col[i] = gf2_apply(a, degree, b.col[i]);
  c = t;
}

// cache of A^(2^j), j < 64, for the most recent (mask, degree)
static std::mutex g_pow_mu;
static uint64_t g_pow_mask = 0;
static int g_pow_degree = 0;
static Gf2Mat g_pow[64];

void gf2_step_powers(uint64_t mask, int degree, Gf2Mat out[64]) {
  std::lock_guard<std::mutex> lk(g_pow_mu);
  if (g_pow_mask != mask || g_pow_degree != degree) {
    gf2_step_matrix(mask, degree, g_pow[0]);
    for (int j = 1; j < 64; ++j) gf2_mul(g_pow[j - 1], g_pow[j - 1], degree, g_pow[j]);
    g_pow_mask = mask;
    g_pow_degree = degree;
  }
  memcpy(out, g_pow, sizeof(g_pow));
}

void gf2_step_power(uint64_t mask, int degree, uint64_t steps, Gf2Mat& out) {
  Gf2Mat pw[64];
  gf2_step_powers(mask, degree, pw);
  memset(&out, 0, sizeof(out));
  for (int i = 0; i < degree; ++i) out.col[i] = 1ull << i;  // identity
  for (int j = 0; j < 64; ++j)
    if ((steps >> j) & 1) gf2_mul(pw[j], out, degree, out);
}

uint64_t lfsr_jump(uint64_t mask, int degree, uint64_t state, uint64_t steps) {
  Gf2Mat pw[64];
  gf2_step_powers(mask, degree, pw);
  for (int j = 0; j < 64; ++j)
    if ((steps >> j) & 1) state = gf2_apply(pw[j], degree, state);
  return state;
}

}  // namespace wf

extern "C" {

int wf_abi_version(void) { return WF_ABI_VERSION; }

const char* wf_last_error_string(void) { return wf::err_buf(); }

int wf_device_check(void) { return wf::require_device(); }

uint64_t wf_lfsr_mask(int degree) { return wf::lfsr_mask(degree); }

uint64_t wf_lfsr_jump(uint64_t mask, int degree, uint64_t state, uint64_t steps) {
  if (degree < 1 || degree > 64) return 0;
  return wf::lfsr_jump(mask, degree, state, steps);
}

}  // extern "C"
