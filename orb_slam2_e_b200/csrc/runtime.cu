// Process-wide runtime bits of liborbfront.so: thread-local error text, device selection,
// launch accounting.
#include <cstring>

#include "common.cuh"

namespace orbfront {

static thread_local char t_err[512] = "";
std::atomic<long long> g_launches{0};

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
  return code;
}

int use_device(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return set_error(ORB_E_NO_DEVICE, "no CUDA device available (%s); liborbfront has no CPU fallback",
                     e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  }
  if (device < 0 || device >= n) return set_error(ORB_E_INVALID, "device %d out of range (0..%d)", device, n - 1);
  e = cudaSetDevice(device);
  if (e != cudaSuccess) return set_error(ORB_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  return ORB_OK;
}

namespace {
constexpr int kMaxScratchDevices = 16;
struct ThreadScratch {
  CallScratch s[kMaxScratchDevices];
  ~ThreadScratch() {  // thread exit; at process teardown the context may be gone already: ignore errors
    for (CallScratch& c : s) {
      if (c.dev) cudaFree(c.dev);
      if (c.host) cudaFreeHost(c.host);
      if (c.stream) cudaStreamDestroy(c.stream);
    }
    cudaGetLastError();
  }
};
thread_local ThreadScratch t_scratch;
}  // namespace

int call_scratch(int device, size_t bytes, CallScratch** out) {
  if (device < 0 || device >= kMaxScratchDevices) return set_error(ORB_E_INVALID, "device %d: at most %d devices", device, kMaxScratchDevices);
  CallScratch& c = t_scratch.s[device];
  if (!c.stream) ORB_CUDA_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  if (bytes > c.cap) {
    const size_t want = align_up<size_t>(bytes + bytes / 2, 1 << 16);
    if (c.dev) cudaFree(c.dev);
    if (c.host) cudaFreeHost(c.host);
    c.dev = nullptr; c.host = nullptr; c.cap = 0;
    ORB_CUDA_TRY(cudaMalloc(&c.dev, want));
    ORB_CUDA_TRY(cudaMallocHost(&c.host, want));
    c.cap = want;
  }
  *out = &c;
  return ORB_OK;
}

}  // namespace orbfront

extern "C" {
const char* orb_last_error(void) { return orbfront::t_err; }
int orb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
long long orb_launch_count(void) { return orbfront::g_launches.load(); }
}
