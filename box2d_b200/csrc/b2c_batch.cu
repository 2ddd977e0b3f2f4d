// libb2cuda — batches of independent worlds (b2cuda_batch_*; see the Batch class in b2c_host.cuh and k_batch_step in b2c_step.cu).
// Host side only: the device arena the member worlds allocate from, the device array of their WorldDev records, one launch per step,
// statistics reduced on the device, per-world reset from a snapshot taken at the first step.
#include "b2c_host.cuh"
#include <cstring>
#include <cstdlib>
#include <algorithm>

namespace b2c {

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(B2C_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

cudaError_t launchBatchStep(const WorldDev* worlds, int worldCount, const StepParams& sp, int newEndIndex, int* maxIslandConsDev, int maxIslandConsSeen, cudaStream_t stream);
cudaError_t launchBatchZeroStep(const WorldDev* worlds, int worldCount, int newEndIndex, cudaStream_t stream);
cudaError_t prepareBatchKernels();
cudaError_t launchBatchStats(const WorldDev* worlds, int worldCount, int lastEndIndex, double* out, unsigned int* errOr, cudaStream_t stream);

cudaStream_t batchStream(Batch* b) { return b->stream(); }
void* batchArenaAlloc(Batch* b, size_t bytes) { return b->arenaAlloc(bytes); }
int* batchMappedInt(Batch* b, int** devPtr) { return b->mappedInt(devPtr); }
void batchMarkDirty(Batch* b, int index) { b->markDirty(index); }

constexpr size_t kArenaChunk = 256ull << 20;

Batch::Batch(const B2CWorldDef& def, int device, int worldCount) : device_(device), def_(def) {
  worlds_.resize(worldCount, nullptr);
}

int Batch::init() {
  CK(cudaSetDevice(device_));
  CK(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
  for (auto& e : timer_) CK(cudaEventCreate(&e));
  CK(cudaHostAlloc((void**)&mappedInt_, sizeof(int), cudaHostAllocMapped));
  CK(cudaHostGetDevicePointer((void**)&mappedIntDev_, mappedInt_, 0));
  CK(prepareBatchKernels());
  CK(cudaHostAlloc((void**)&maxCons_, sizeof(int), cudaHostAllocMapped));
  *maxCons_ = 0;
  CK(cudaHostGetDevicePointer((void**)&maxConsDev_, maxCons_, 0));
  CK(cudaMalloc((void**)&dStats_, sizeof(double) * 10));
  CK(cudaMemsetAsync(dStats_, 0, sizeof(double) * 10, stream_));
  CK(cudaHostAlloc((void**)&hStats_, sizeof(double) * 8, cudaHostAllocDefault));
  for (size_t i = 0; i < worlds_.size(); i++) {
    worlds_[i] = new (std::nothrow) World(def_, device_, this, (int)i);
    if (!worlds_[i]) return fail(B2C_E_INVALID, "out of host memory");
  }
  return B2C_OK;
}

Batch::~Batch() {
  if (device_ >= 0) cudaSetDevice(device_);
  if (stream_) cudaStreamSynchronize(stream_);
  for (World* w : worlds_) delete w;               // members free no device memory: it is the arena's
  for (void* c : chunks_) cudaFree(c);
  for (void* c : snapChunks_) cudaFree(c);
  for (auto& c : graphs_) if (c.exec) cudaGraphExecDestroy(c.exec);
  if (dWorlds_) cudaFree(dWorlds_);
  if (dStats_) cudaFree(dStats_);
  if (hStats_) cudaFreeHost(hStats_);
  if (mappedInt_) cudaFreeHost(mappedInt_);
  if (maxCons_) cudaFreeHost(maxCons_);
  for (auto& e : timer_) if (e) cudaEventDestroy(e);
  if (stream_) cudaStreamDestroy(stream_);
}

// Two arena chunks can be neighbours in the address space: a run of adjacent blocks must still end where a chunk begins (one copy, one allocation).
bool Batch::isChunkBase(const void* p) const { return std::find(chunks_.begin(), chunks_.end(), p) != chunks_.end(); }
void* Batch::arenaAlloc(size_t bytes) {
  bytes = (bytes + 255) & ~(size_t)255;
  if (bytes > chunkLeft_) {
    size_t sz = std::max(bytes, kArenaChunk);
    void* p = nullptr;
    if (cudaMalloc(&p, sz) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    cudaMemsetAsync(p, 0, sz, stream_);
    chunks_.push_back(p); chunkPtr_ = (char*)p; chunkLeft_ = sz;
  }
  void* r = chunkPtr_;
  chunkPtr_ += bytes; chunkLeft_ -= bytes;
  return r;
}

// First step: every world that has not been uploaded yet goes to the device, the WorldDev records are gathered into one device array, and
// the state of every world (device arrays + the host mirror) is kept for resetWorld.
int Batch::ensureUploaded() {
  if (uploaded_) return B2C_OK;
  CK(cudaSetDevice(device_));
  const int n = (int)worlds_.size();
  for (int i = 0; i < n; i++) {
    World& w = *worlds_[i];
    if (!w.onDevice_) { int rc = w.upload(); if (rc) return fail(rc, "world " + std::to_string(i) + ": " + w.error); }
  }
  std::vector<WorldDev> host(n);
  for (int i = 0; i < n; i++) { host[i] = worlds_[i]->d_; worlds_[i]->devDirty_ = false; }
  dirty_.clear();
  CK(cudaMalloc((void**)&dWorlds_, sizeof(WorldDev) * (size_t)std::max(n, 1)));
  CK(cudaMemcpyAsync(dWorlds_, host.data(), sizeof(WorldDev) * (size_t)n, cudaMemcpyHostToDevice, stream_));
  CK(cudaStreamSynchronize(stream_));
  // snapshot: one device block per world (its arena blocks back to back), chunked like the arena
  snaps_.resize(n);
  char* sp = nullptr; size_t left = 0;
  for (int i = 0; i < n; i++) {
    World& w = *worlds_[i];
    size_t total = 0;
    for (auto& b : w.arenaBlocks_) total += (b.second + 255) & ~(size_t)255;
    if (total > left) {
      size_t sz = std::max(total, kArenaChunk);
      void* p = nullptr;
      if (cudaMalloc(&p, sz) != cudaSuccess) { cudaGetLastError(); return fail(B2C_E_CUDA, "cudaMalloc failed for the batch reset snapshot"); }
      snapChunks_.push_back(p); sp = (char*)p; left = sz;
    }
    Snap& s = snaps_[i];
    s.dev = sp; s.bytes = total;
    size_t off = 0;
    // consecutive arena blocks are adjacent in memory (bump allocation) unless a chunk boundary fell between them: copy runs
    for (size_t k = 0; k < w.arenaBlocks_.size();) {
      char* runBegin = (char*)w.arenaBlocks_[k].first; size_t run = 0; size_t j = k;
      while (j < w.arenaBlocks_.size() && (char*)w.arenaBlocks_[j].first == runBegin + run && (j == k || !isChunkBase(w.arenaBlocks_[j].first))) { run += (w.arenaBlocks_[j].second + 255) & ~(size_t)255; j++; }
      CK(cudaMemcpyAsync(s.dev + off, runBegin, run, cudaMemcpyDeviceToDevice, stream_));
      off += run; k = j;
    }
    sp += total; left -= total;
    w.saveHostSnap(s.host);
  }
  uploaded_ = true;
  return B2C_OK;
}

int Batch::resetWorld(int index) {
  if (index < 0 || index >= (int)worlds_.size()) return fail(B2C_E_INVALID, "bad world index");
  if (!uploaded_) return B2C_OK;                   // never stepped: nothing to undo
  CK(cudaSetDevice(device_));
  World& w = *worlds_[index];
  Snap& s = snaps_[index];
  size_t off = 0;
  for (size_t k = 0; k < w.arenaBlocks_.size();) {
    char* runBegin = (char*)w.arenaBlocks_[k].first; size_t run = 0; size_t j = k;
    while (j < w.arenaBlocks_.size() && (char*)w.arenaBlocks_[j].first == runBegin + run && (j == k || !isChunkBase(w.arenaBlocks_[j].first))) { run += (w.arenaBlocks_[j].second + 255) & ~(size_t)255; j++; }
    CK(cudaMemcpyAsync(runBegin, s.dev + off, run, cudaMemcpyDeviceToDevice, stream_));
    off += run; k = j;
  }
  w.restoreHostSnap(s.host);
  w.scalarsStale_ = true;
  w.devDirty_ = false; w.markDeviceDirty();
  return B2C_OK;
}

int Batch::step(float timeStep, int subStepCount) {
  if (def_.enableSleep) return fail(B2C_E_UNSUPPORTED, "enableSleep=true is not supported for batches; create the batch with enableSleep=0");
  { int rc = ensureUploaded(); if (rc) return rc; }
  CK(cudaSetDevice(device_));
  const int n = (int)worlds_.size();
  if (n == 0) return B2C_OK;
  for (int i : dirty_) {                           // worlds a creation / destruction / setter call touched since the last step
    World& w = *worlds_[i];
    CK(cudaMemcpyAsync(dWorlds_ + i, &w.d_, sizeof(WorldDev), cudaMemcpyHostToDevice, stream_));
    w.devDirty_ = false;
  }
  if (!dirty_.empty()) { CK(cudaStreamSynchronize(stream_)); dirty_.clear(); }   // pageable sources
  const int lastEnd = endIndex_;
  endIndex_ = 1 - endIndex_;
  if (timeStep == 0.0f) {                          // api/World.cs:44-51
    CK(launchBatchZeroStep(dWorlds_, n, endIndex_, stream_));
  } else {
    StepParams sp;
    memset(&sp, 0, sizeof(sp));
    sp.dt = timeStep; sp.subStepCount = subStepCount > 1 ? subStepCount : 1;
    if (timeStep > 0.0f) { sp.inv_dt = 1.0f / timeStep; sp.h = timeStep / sp.subStepCount; sp.inv_h = sp.subStepCount * sp.inv_dt; }
    const World& w0 = *worlds_[0];                 // tunables: the batch definition's, as every member was created with it
    float hz = netMin(w0.contactHertz, 0.125f * sp.inv_h);     // api/World.cs:63-83
    sp.contactSoftness = makeSoft(hz, w0.contactDampingRatio, sp.h);
    sp.staticSoftness = makeSoft(2.0f * hz, w0.contactDampingRatio, sp.h);
    sp.restitutionThreshold = w0.restitutionThreshold; sp.maxLinearSpeed = w0.maxLinearSpeed; sp.contactSpeed = w0.contactSpeed; sp.contactHertz = w0.contactHertz;
    sp.contactDampingRatio = w0.contactDampingRatio; sp.hitEventThreshold = w0.hitEventThreshold; sp.gravity = w0.gravity;
    sp.enableWarmStarting = w0.enableWarmStarting; sp.enableContactSoftening = w0.enableContactSoftening; sp.enableContinuous = w0.enableContinuous;
    sp.enableSpeculative = w0.enableSpeculative; sp.enableSleep = 0; sp.endIndex = lastEnd;
    sp.clusterThreads = 0;
    const int maxCons = *(volatile int*)maxCons_;   // the maximum may lag the running step: either CTA width is correct
    // The launches of a step are replayed as ONE CUDA-graph launch (captured per distinct step parameters / end-buffer parity / CTA
    // width): with several ranks driving GPUs from one host the three kernels then run back to back whatever the host's launch latency is.
    static const bool useGraph = !(getenv("B2CUDA_BATCH_GRAPH") && atoi(getenv("B2CUDA_BATCH_GRAPH")) == 0);
    if (useGraph) {
      StepGraphB* g = nullptr;
      const int narrow = maxCons > 0 && maxCons <= 160;
      for (auto& c : graphs_) if (c.endIndex == endIndex_ && c.narrow == narrow && c.n == n && memcmp(&c.sp, &sp, sizeof(sp)) == 0) { g = &c; break; }
      if (!g) {
        if (graphs_.size() >= 8) { for (auto& c : graphs_) cudaGraphExecDestroy(c.exec); graphs_.clear(); }
        cudaGraph_t graph = nullptr;
        CK(cudaStreamBeginCapture(stream_, cudaStreamCaptureModeThreadLocal));
        cudaError_t le = launchBatchStep(dWorlds_, n, sp, endIndex_, maxConsDev_, maxCons, stream_);
        cudaError_t ce = cudaStreamEndCapture(stream_, &graph);
        if (le != cudaSuccess || ce != cudaSuccess) { if (graph) cudaGraphDestroy(graph); cudaGetLastError(); return fail(B2C_E_CUDA, std::string("batch step capture: ") + cudaGetErrorString(le != cudaSuccess ? le : ce)); }
        StepGraphB c; c.sp = sp; c.endIndex = endIndex_; c.narrow = narrow; c.n = n; c.exec = nullptr;
        cudaError_t ie = cudaGraphInstantiate(&c.exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ie != cudaSuccess) return fail(B2C_E_CUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ie));
        graphs_.push_back(c); g = &graphs_.back();
      }
      CK(cudaGraphLaunch(g->exec, stream_));
    } else CK(launchBatchStep(dWorlds_, n, sp, endIndex_, maxConsDev_, maxCons, stream_));
  }
  const bool zero = timeStep == 0.0f;
  for (World* w : worlds_) {                       // per-world host bookkeeping of World::step
    w->lastEndIndex_ = lastEnd; w->endIndex_ = endIndex_; w->lastStepDtZero_ = zero;
    w->eventsFetched_ = false; w->bodiesFetched_ = false; w->movesFetched_ = false; w->scalarsStale_ = true; w->stepCount_++;
  }
  stepCount_++;
  return B2C_OK;
}

int Batch::synchronize() {
  CK(cudaSetDevice(device_));
  if (stream_) CK(cudaStreamSynchronize(stream_));
  return B2C_OK;
}

int Batch::stats(B2CBatchStats* out, double* deviceDst) {
  { int rc = ensureUploaded(); if (rc) return rc; }
  CK(cudaSetDevice(device_));
  CK(launchBatchStats(dWorlds_, (int)worlds_.size(), 1 - endIndex_, dStats_, (unsigned int*)(dStats_ + 8), stream_));
  if (deviceDst) CK(cudaMemcpyAsync(deviceDst, dStats_, sizeof(double) * 8, cudaMemcpyDeviceToDevice, stream_));
  CK(cudaMemcpyAsync(hStats_, dStats_, sizeof(double) * 8, cudaMemcpyDeviceToHost, stream_));
  CK(cudaStreamSynchronize(stream_));
  if (out) { out->awakeBodies = hStats_[0]; out->sumV2 = hStats_[1]; out->touchingContacts = hStats_[2]; out->beginEvents = hStats_[3];
             out->endEvents = hStats_[4]; out->maxSpeed = hStats_[5]; out->minY = hStats_[6]; out->errorFlags = hStats_[7]; }
  return B2C_OK;
}

int Batch::timerStart() {
  { int rc = ensureUploaded(); if (rc) return rc; }
  CK(cudaSetDevice(device_));
  CK(cudaEventRecord(timer_[0], stream_));
  return B2C_OK;
}
int Batch::timerStop(float* ms) {
  *ms = 0;
  CK(cudaSetDevice(device_));
  CK(cudaEventRecord(timer_[1], stream_));
  CK(cudaEventSynchronize(timer_[1]));
  CK(cudaEventElapsedTime(ms, timer_[0], timer_[1]));
  return B2C_OK;
}

}  // namespace b2c
