// api.cu -- the extern "C" surface declared in include/msgat_sm100.h.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "compat.cuh"
#include "kernels.h"

static thread_local char g_err[512] = "";

void msgat_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ---------------------------------------------------------------------------------------
// launch accounting: every kernel launch of this library is counted (bench.py reports the
// count as gpu_launches); with msgat_profile(1) each launch is bracketed by CUDA events on
// its own stream so that bench.py can attribute device time to kernels without a profiler.
#ifndef MSGAT_EMU
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_prof{0};
struct ProfRec { std::string name; cudaEvent_t e0, e1; };
static std::vector<ProfRec> g_recs;
static std::mutex g_prof_mu;
static thread_local char g_tag[48] = "";
void msgat_prof_tag(const char* fmt, int a, int b) { snprintf(g_tag, sizeof(g_tag), fmt, a, b); }
int msgat_prof_begin(const char* name, cudaStream_t s) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (!g_prof.load(std::memory_order_relaxed)) return -1;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfRec r;
  r.name = std::string(name) + g_tag;
  cudaEventCreate(&r.e0); cudaEventCreate(&r.e1);
  cudaEventRecord(r.e0, s);
  g_recs.push_back(r);
  return (int)g_recs.size() - 1;
}
int msgat_current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev;
}
int msgat_sm_count() {
  static std::atomic<int> cache[64];
  const int dev = msgat_current_device() & 63;
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n <= 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}
void msgat_prof_end(int slot, cudaStream_t s) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  cudaEventRecord(g_recs[slot].e1, s);
}

// ---- a side stream per device: independent small kernels of one operator run next to its big ones (fork / join with events, so
// the pattern is legal under CUDA graph capture: the side stream joins the capture at the fork and leaves it at the join).  All
// side work is joined back into the caller's stream before the operator returns: callers see one stream, as the C ABI says.
struct SideCtx { cudaStream_t s; cudaEvent_t fork, join; bool ok; };
static SideCtx g_side[64];
static std::mutex g_side_mu;
static SideCtx* side_ctx() {
  static const bool off = getenv("MSGAT_NO_FORK") != nullptr;   // measurement switch: everything on the caller's stream
  if (off) return nullptr;
  const int dev = msgat_current_device() & 63;
  std::lock_guard<std::mutex> lk(g_side_mu);
  SideCtx& c = g_side[dev];
  if (!c.ok) {
    if (cudaStreamCreateWithFlags(&c.s, cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c.join, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    c.ok = true;
  }
  return &c;
}
void* msgat_fork(void* stream) {
  SideCtx* c = side_ctx();
  if (!c) return stream;
  if (cudaEventRecord(c->fork, (cudaStream_t)stream) != cudaSuccess || cudaStreamWaitEvent(c->s, c->fork, 0) != cudaSuccess) {
    cudaGetLastError();
    return stream;
  }
  return (void*)c->s;
}
void msgat_join(void* stream, void* side) {
  if (side == stream) return;
  SideCtx* c = side_ctx();
  cudaEventRecord(c->join, (cudaStream_t)side);
  cudaStreamWaitEvent((cudaStream_t)stream, c->join, 0);
}
#endif

static int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    msgat_set_error("%s: %s", what, cudaGetErrorString(e));
    return -100;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------
// adjacency: dense [N,N] -> CSR + CSC (data_loader.py:59-66 produces the dense matrix;
// msgat.py:190 stores it).  Built once per adjacency by the Python wrapper.
MSGAT_KERNEL k_adj_count(const float* adj, int N, int32_t* row_nnz, int32_t* col_nnz) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  int rc = 0, cc = 0;
  for (int j = 0; j < N; ++j) {
    rc += adj[(int64_t)r * N + j] != 0.f;
    cc += adj[(int64_t)j * N + r] != 0.f;
  }
  row_nnz[r] = rc;
  col_nnz[r] = cc;
}

MSGAT_KERNEL k_adj_fill_rows(const float* adj, int N, const int32_t* row_ptr, int32_t* col_idx, float* val) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  int e = row_ptr[r];
  for (int j = 0; j < N; ++j) {
    const float v = adj[(int64_t)r * N + j];
    if (v != 0.f) { col_idx[e] = j; val[e] = v; ++e; }
  }
}

MSGAT_KERNEL k_adj_fill_cols(const float* adj, int N, const int32_t* row_ptr, const int32_t* col_idx,
                             const int32_t* csc_ptr, int32_t* csc_row, int32_t* csc_eid) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  int k = csc_ptr[c];
  for (int r = 0; r < N; ++r) {
    if (adj[(int64_t)r * N + c] == 0.f) continue;
    // position of column c inside CSR row r (columns ascending)
    int e = row_ptr[r];
    while (col_idx[e] != c) ++e;
    csc_row[k] = r;
    csc_eid[k] = e;
    ++k;
  }
}

extern "C" {

const char* msgat_version(void) { return "msgat_sm100 0.1 (sm_100a)"; }
const char* msgat_last_error(void) { return g_err; }

long long msgat_launch_count(void) {
#ifdef MSGAT_EMU
  return 0;
#else
  return g_launches.load();
#endif
}

int msgat_profile(int enable) {
#ifndef MSGAT_EMU
  g_prof.store(enable ? 1 : 0);
#endif
  return 0;
}

/* Synchronises the recorded events and writes "kernel launches total_ms\n" lines. */
int msgat_profile_report(char* buf, long long cap) {
#ifdef MSGAT_EMU
  if (cap > 0) buf[0] = 0;
  return 0;
#else
  std::lock_guard<std::mutex> lk(g_prof_mu);
  std::map<std::string, std::pair<long long, double>> agg;
  for (auto& r : g_recs) {
    cudaEventSynchronize(r.e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r.e0, r.e1);
    auto& a = agg[r.name];
    a.first += 1; a.second += ms;
    cudaEventDestroy(r.e0); cudaEventDestroy(r.e1);
  }
  g_recs.clear();
  std::string out;
  for (auto& kv : agg) {
    char line[256];
    snprintf(line, sizeof(line), "%s %lld %.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
    out += line;
  }
  if ((long long)out.size() + 1 > cap) { msgat_set_error("profile report needs %zu bytes", out.size() + 1); return -50; }
  memcpy(buf, out.c_str(), out.size() + 1);
  return 0;
#endif
}

#ifdef MSGAT_EMU
/* the host-fiber debug build has no tensor cores: tc_*.cu is not part of it */
int msgat_tc_probe(int, const float*, const float*, float*, int, int, int, int, int, void*) {
  msgat_set_error("tcgen05 kernels need the sm_100a build");
  return -200;
}
#endif

int msgat_device_check(void) {
#ifdef MSGAT_EMU
  return 0;
#else
  int dev = 0;
  cudaDeviceProp prop;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
    msgat_set_error("no CUDA device");
    return -1;
  }
  if (prop.major != 10) {
    msgat_set_error("libmsgat_sm100 needs compute capability 10.x (B200), found %d.%d", prop.major, prop.minor);
    return -2;
  }
  return 0;
#endif
}

int msgat_adj_count(const float* adj, int32_t N, int32_t* row_nnz, int32_t* col_nnz, void* stream) {
  MSGAT_LAUNCH(k_adj_count, dim3((N + 127) / 128), dim3(128), 0, stream, adj, N, row_nnz, col_nnz);
  return check_launch("adj_count");
}

int msgat_adj_fill(const float* adj, int32_t N, const int32_t* row_ptr, const int32_t* csc_ptr, int32_t* col_idx,
                   float* val, int32_t* csc_row, int32_t* csc_eid, void* stream) {
  MSGAT_LAUNCH(k_adj_fill_rows, dim3((N + 127) / 128), dim3(128), 0, stream, adj, N, row_ptr, col_idx, val);
  MSGAT_LAUNCH(k_adj_fill_cols, dim3((N + 127) / 128), dim3(128), 0, stream, adj, N, row_ptr, (const int32_t*)col_idx,
               csc_ptr, csc_row, csc_eid);
  return check_launch("adj_fill");
}

int msgat_meam_sizes(const msgat_meam_desc* d, int32_t nnz, int64_t* saved_floats, int64_t* scratch_floats) {
  Plan p;
  int rc = make_plan(d, nnz, &p);
  if (rc) { msgat_set_error("unsupported MEAM descriptor (code %d)", rc); return rc; }
  int64_t bs = meam_backward_scratch(p);
  if (bs < 0) { msgat_set_error("unsupported MEAM descriptor for backward (code %lld)", (long long)bs); return (int)bs; }
  *saved_floats = p.saved_total;
  *scratch_floats = p.fwd_scratch > bs ? p.fwd_scratch : bs;
  return 0;
}

// floats between the scratch buffers of consecutive components: what msgat_meam_sizes reported
static int set_scratch_stride(Plan& p) {
  const int64_t bs = meam_backward_scratch(p);
  if (bs < 0) { msgat_set_error("unsupported MEAM descriptor for backward (code %lld)", (long long)bs); return (int)bs; }
  p.scratch_sm = p.fwd_scratch > bs ? p.fwd_scratch : bs;
  return 0;
}

int msgat_meam_forward(const msgat_meam_desc* d, const msgat_meam_params* prm, const msgat_adj* adj, const float* x,
                       float* out, float* saved, float* scratch, void* stream) {
  Plan p;
  int rc = make_plan(d, adj ? adj->nnz : 0, &p);
  if (rc) { msgat_set_error("unsupported MEAM descriptor (code %d)", rc); return rc; }
  if (p.Os && !adj) { msgat_set_error("spatial branch needs an adjacency"); return -20; }
  if ((rc = set_scratch_stride(p))) return rc;
  msgat_prof_tag("@C%dO%d", p.C, p.O);
  rc = meam_forward_impl(p, prm, adj, x, out, saved, scratch, stream);
  if (rc) return rc;
  return check_launch("meam_forward");
}

int msgat_meam_backward(const msgat_meam_desc* d, const msgat_meam_params* prm, const msgat_adj* adj, const float* x,
                        const float* out, const float* dout, const float* saved, float* scratch, float* dx,
                        const msgat_meam_grads* g, void* stream) {
  Plan p;
  int rc = make_plan(d, adj ? adj->nnz : 0, &p);
  if (rc) { msgat_set_error("unsupported MEAM descriptor (code %d)", rc); return rc; }
  if (p.Os && !adj) { msgat_set_error("spatial branch needs an adjacency"); return -20; }
  if ((rc = set_scratch_stride(p))) return rc;
  msgat_prof_tag("@C%dO%d", p.C, p.O);
  rc = meam_backward_impl(p, prm, adj, x, out, dout, saved, scratch, dx, g, stream);
  if (rc) return rc;
  return check_launch("meam_backward");
}

}  // extern "C"
