// libb2cuda — b2cuda_world_copy_body_move_events_noalloc: the body move events without any allocation.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it). It is a variant of the pack kernel behind
// b2cuda_world_get_body_move_events (k_pack_moves in b2c_step.cu), which packs into a per-world array obtained from cudaMalloc
// and a pinned host buffer obtained from cudaHostAlloc. Here the pack kernel writes into ONE file-scope device array of a fixed
// maximum size and the records are then copied into the caller's buffer: no cudaMalloc, cudaFree, cudaHostAlloc or any other
// allocating routine is called, on any call.
//
// The scratch array
//   * holds B2C_NOALLOC_MAX_MOVE_EVENTS records (65 536 x 24 B = 1.5 MiB); every GPU that loads the library holds it until exit;
//   * is shared by every call, of every world, on that GPU: calls that overlap on the GPU (different streams / host threads)
//     race on it. Each call synchronises its world's stream before it returns, so calls made one after another are safe;
//   * keeps the previous call's contents. Nothing has to start at zero: the kernel overwrites records [0, count) before they
//     are copied out, and records past `count` are never read.
// In host code the array's NAME is a host-side stand-in, not a device address; the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current (scratchFor below).
#include "b2c_host.cuh"
#include <mutex>
#include <string>

namespace b2c {

static __device__ __align__(256) B2CBodyMoveEvent g_moveScratch[B2C_NOALLOC_MAX_MOVE_EVENTS];

constexpr int kNoAllocMaxDevices = 64;
static B2CBodyMoveEvent* g_moveScratchPtr[kNoAllocMaxDevices] = {};
static std::mutex g_scratchMutex;

// Device address of g_moveScratch on `device` (which must be current). Asked once per device, then cached.
static cudaError_t scratchFor(int device, B2CBodyMoveEvent** out) {
  *out = nullptr;
  if (device < 0 || device >= kNoAllocMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_scratchMutex);
  if (!g_moveScratchPtr[device]) {
    B2CBodyMoveEvent* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_moveScratch);
    if (e != cudaSuccess) return e;
    if (!p) return cudaErrorInvalidSymbol;
    g_moveScratchPtr[device] = p;
  }
  *out = g_moveScratchPtr[device];
  return cudaSuccess;
}

// BodyMoveEvent per awake body in awake-set order (src/Solver.cs:432-438) — the same records k_pack_moves writes
__global__ void __launch_bounds__(256) k_pack_moves_scratch(WorldDev d, B2CBodyMoveEvent* scratch, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    int bodyId = d.simBodyId[i];
    float4 xf = d.bodyXf[bodyId];
    B2CBodyMoveEvent ev; ev.transform.p = {xf.x, xf.y}; ev.transform.q = {xf.z, xf.w}; ev.bodyId = bodyId; ev.fellAsleep = 0;
    scratch[i] = ev;
  }
}

static int noAllocMoveEvents(World& w, B2CBodyMoveEvent* dst, int capacity, int* count) {
  *count = 0;
  const World::NoAllocView v = w.noAllocView();
  auto fail = [&](int code, const std::string& msg) { w.error = msg; return code; };
  if (!v.onDevice) return fail(B2C_E_INVALID, "copy_body_move_events_noalloc before the first step");
  const int n = v.dev->awakeCount;
  if (n > B2C_NOALLOC_MAX_MOVE_EVENTS)
    return fail(B2C_E_CAPACITY, "copy_body_move_events_noalloc: " + std::to_string(n) + " awake bodies exceed the fixed scratch of " +
                                std::to_string(B2C_NOALLOC_MAX_MOVE_EVENTS) + " records (B2C_NOALLOC_MAX_MOVE_EVENTS); use b2cuda_world_get_body_move_events");
  const int reported = v.lastStepDtZero ? 0 : n;          // api/World.cs:38,44-51: a dt = 0 step clears the move events
  if (reported > capacity) return fail(B2C_E_INVALID, "copy_body_move_events_noalloc: destination holds " + std::to_string(capacity) + " records, " + std::to_string(reported) + " needed");
  if (reported == 0) return B2C_OK;
  cudaError_t e = cudaSetDevice(v.device);
  if (e != cudaSuccess) return fail(B2C_E_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
  B2CBodyMoveEvent* scratch = nullptr;
  e = scratchFor(v.device, &scratch);
  if (e != cudaSuccess) return fail(B2C_E_CUDA, std::string("cudaGetSymbolAddress(g_moveScratch): ") + cudaGetErrorString(e));
  int blocks = (n + 255) / 256;
  if (blocks > 1184) blocks = 1184;                       // 148 SMs x 8
  k_pack_moves_scratch<<<blocks, 256, 0, v.stream>>>(*v.dev, scratch, n);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(B2C_E_CUDA, std::string("k_pack_moves_scratch: ") + cudaGetErrorString(e));
  e = cudaMemcpyAsync(dst, scratch, sizeof(B2CBodyMoveEvent) * (size_t)n, cudaMemcpyDefault, v.stream);   // dst: host (pageable or pinned) or device memory
  if (e != cudaSuccess) return fail(B2C_E_CUDA, std::string("cudaMemcpyAsync: ") + cudaGetErrorString(e));
  e = cudaStreamSynchronize(v.stream);
  if (e != cudaSuccess) return fail(B2C_E_CUDA, std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e));
  *count = n;
  return B2C_OK;
}

}  // namespace b2c

struct B2CWorld { b2c::World* w; };                       // as in b2c_capi.cu
void b2cSetLastError(const std::string& msg);

extern "C" int32_t b2cuda_world_copy_body_move_events_noalloc(B2CWorld* h, B2CBodyMoveEvent* dst, int32_t capacity, int32_t* count) {
  if (!h || !h->w) { b2cSetLastError("null world handle"); return B2C_E_INVALID; }
  if (!dst || !count || capacity < 0) { b2cSetLastError("copy_body_move_events_noalloc: null destination / count or negative capacity"); return B2C_E_INVALID; }
  int n = 0;
  int rc = b2c::noAllocMoveEvents(*h->w, dst, capacity, &n);
  *count = n;
  if (rc != B2C_OK) b2cSetLastError(h->w->error);
  return rc;
}
