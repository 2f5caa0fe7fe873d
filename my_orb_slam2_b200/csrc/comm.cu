// comm.cu — the multi-GPU step of the brute-force map matching (BASELINE.json config 5) behind the
// C ABI: one process (or thread) per GPU, the MAP descriptors sharded, the queries replicated; every rank computes
// its local top-2 keys (k_match.cu) and the ranks exchange them so that every rank ends with the GLOBAL top-2.
//
// Two exchange paths, same results:
//  * peer memory (default when every pair of ranks can map each other's memory: NVLink / NVSwitch): the kernel that
//    merges a rank's slices stores its (best, second) keys straight into EVERY rank's exchange buffer with plain
//    stores over NVLink (the buffers are mapped once, at orbfe_comm_create, through CUDA IPC handles), the last CTA
//    publishes an epoch flag on every peer (fence.sys + st.release.sys), and the kernel that consumes the keys spins on
//    its own flags (ld.acquire.sys, bounded by a timeout) and takes the global top-2.  No collective call, no
//    staging copy: brute force -> merge+push -> wait+merge, three launches per call.
//  * NCCL (fallback, and the comparison the bench reports): local merge into the rank's slot of the gather buffer,
//    one in-place ncclAllGather, the same consuming merge.  libnccl.so.2 is dlopen'ed (the library links only
//    libcudart); it is also what bootstraps the communicator (unique id -> ncclCommInitRank) and carries the IPC
//    handles at set-up.
//
// Buffer reuse: the exchange buffer is double-buffered by call parity.  Rank X can only push call k+2 after its own
// consume of call k+1 returned, which waited for every rank's push of k+1, which those ranks issued (stream order)
// after their consume of call k: so nobody overwrites keys that are still being read.  All calls of one communicator
// must be issued in the same order on every rank (collective semantics) and are serialised on the matcher's stream.
#include <dlfcn.h>
#include <unistd.h>

#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "matcher_ctx.cuh"

namespace orbfe {

// ---- the slice of the NCCL API this file uses (declared here: no build-time dependency on nccl.h) ----
struct NcclUniqueId { char internal[128]; };
typedef void* NcclComm;
constexpr int kNcclUint8 = 1, kNcclUint64 = 5;
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*GetVersion)(int*) = nullptr;
};

static NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.lib ? &api : nullptr;
  tried = true;
  const char* names[] = {getenv("ORBFE_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    if (!n || !n[0]) continue;
    api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) {
    set_error("comm: cannot dlopen libnccl.so.2 (%s); set ORBFE_NCCL_LIB to its path", dlerror());
    return nullptr;
  }
  bool ok = true;
  auto sym = [&](const char* name) {
    void* p = dlsym(api.lib, name);
    if (!p) { set_error("comm: %s missing from the NCCL library", name); ok = false; }
    return p;
  };
  api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
  api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
  api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
  api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
  api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
  if (!ok) { dlclose(api.lib); api.lib = nullptr; return nullptr; }
  return &api;
}

#define ORBFE_NCCL(api, call)                                                                            \
  do {                                                                                                   \
    int r__ = (call);                                                                                    \
    if (r__ != 0) {                                                                                      \
      orbfe::set_error("%s:%d: %s -> NCCL error %d (%s)", __FILE__, __LINE__, #call, r__,                \
                       (api)->GetErrorString ? (api)->GetErrorString(r__) : "?");                        \
      return ORBFE_ERR_CUDA;                                                                             \
    }                                                                                                    \
  } while (0)

constexpr int kMaxRanks = 64;
// a peer that never arrives: the consuming kernel gives up after ORBFE_COMM_TIMEOUT_S seconds (default 30) and reports
static unsigned long long wait_timeout_ns() {
  const char* e = getenv("ORBFE_COMM_TIMEOUT_S");
  const double sec = e && e[0] ? atof(e) : 30.0;
  return (unsigned long long)((sec > 0.01 ? sec : 0.01) * 1e9);
}

// what a rank tells the others at set-up
struct PeerCard {
  cudaIpcMemHandle_t xbuf;   // 64 bytes
  unsigned long long ptr;    // the buffer's address in the owner's process (same-process ranks use it directly)
  long long pid;
  long long host_hash;
  int device;
  int pad;
};

struct PeerTable {           // kernel argument: every rank's exchange buffer as mapped HERE
  unsigned long long* keys[kMaxRanks];
  unsigned int* flags[kMaxRanks];
};

struct Comm {
  NcclApi* api = nullptr;
  NcclComm nccl = nullptr;
  int nranks = 1, rank = 0, device = 0, max_queries = 0;
  bool p2p = false;
  // exchange buffer of THIS rank: [parity 2][src rank][best|second][max_queries] u64, then [parity 2][src rank] u32 flags,
  // then a done counter and an error word
  uint8_t* xbuf = nullptr;
  size_t keys_bytes = 0;
  std::vector<void*> opened;  // cudaIpcOpenMemHandle mappings to close
  PeerTable peers;
  unsigned int epoch = 0;
  unsigned int* h_err = nullptr;  // pinned: the consuming kernel reports a timeout here
  // last call's timing events (bench breakdown): kernel | exchange+merge
  cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
  bool timing = false;
};

static inline size_t keys_off(const Comm* c, int parity, int src, int which) {
  return (((size_t)parity * c->nranks + src) * 2 + which) * (size_t)c->max_queries;
}
static inline unsigned int* flags_of(uint8_t* xbuf, size_t keys_bytes) {
  return reinterpret_cast<unsigned int*>(xbuf + keys_bytes);
}

// ---- kernels ----
__device__ __forceinline__ void top2_insert(unsigned long long x, unsigned long long y, unsigned long long& b1,
                                            unsigned long long& b2) {
  if (x < b1) { b2 = b1; b1 = x; } else if (x < b2) b2 = x;
  if (y < b2) b2 = y;  // y >= x: it can only be a second best
}

// merge this rank's slices (arrays [slice][nq]) and store the pair for query qi into EVERY rank's exchange buffer
// (peer stores over NVLink; own buffer included); the last CTA to finish publishes `epoch` in flag[parity][rank] of
// every rank.  done_counter is this rank's own (device-scope) CTA counter, left at zero.
__global__ void __launch_bounds__(256) merge_push_kernel(const unsigned long long* __restrict__ part_best,
                                                         const unsigned long long* __restrict__ part_second, int slices,
                                                         int nq, PeerTable peers, int nranks, int rank, size_t slot_best,
                                                         size_t slot_second, int flag_index, unsigned int epoch,
                                                         unsigned int* done_counter) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi < nq) {
    unsigned long long b1 = ~0ull, b2 = ~0ull;
    for (int p = 0; p < slices; ++p) top2_insert(part_best[(size_t)p * nq + qi], part_second[(size_t)p * nq + qi], b1, b2);
    for (int r = 0; r < nranks; ++r) {
      const int dst = (rank + r) % nranks;  // ranks start on different links
      peers.keys[dst][slot_best + qi] = b1;
      peers.keys[dst][slot_second + qi] = b2;
    }
  }
  __threadfence_system();  // this thread's peer stores are ordered before what follows, at system scope
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done_counter, 1u);
    if (prev == gridDim.x - 1) {  // every CTA of this launch has fenced its stores
      __threadfence_system();
      *done_counter = 0u;
      for (int r = 0; r < nranks; ++r) {
        unsigned int* f = peers.flags[(rank + r) % nranks] + flag_index;
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(epoch) : "memory");
      }
    }
  }
}

// wait until every rank has published `epoch` in this rank's flags (P2P path; nranks_wait = 0 skips the wait: the
// NCCL all-gather is ordered by the stream), then take the top-2 of the nranks (best, second) pairs per query.
__global__ void __launch_bounds__(256) wait_merge_kernel(const unsigned long long* __restrict__ keys, size_t parity_base,
                                                         int max_queries, int nranks, int nq,
                                                         const unsigned int* flags, int nranks_wait, unsigned int epoch,
                                                         unsigned int* err, unsigned long long timeout_ns,
                                                         unsigned long long* __restrict__ best,
                                                         unsigned long long* __restrict__ second) {
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if ((int)threadIdx.x < nranks_wait) {
    const unsigned int* f = flags + threadIdx.x;
    unsigned long long t0 = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      unsigned int v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
      if ((int)(v - epoch) >= 0) break;
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > timeout_ns) { s_ok = 0; break; }
      __nanosleep(200);
    }
  }
  __syncthreads();
  if (!s_ok) {
    if (threadIdx.x == 0) *err = 1u;  // pinned host word: orbfe_comm_status reports it
    return;
  }
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  unsigned long long b1 = ~0ull, b2 = ~0ull;
  for (int r = 0; r < nranks; ++r) {
    const unsigned long long* k = keys + parity_base + (size_t)r * 2 * max_queries;
    top2_insert(__ldcg(k + qi), __ldcg(k + max_queries + qi), b1, b2);  // L2 reads: the peers wrote these lines
  }
  best[qi] = b1;
  second[qi] = b2;
}

static void comm_free(Comm* c) {
  if (!c) return;
  DeviceGuard guard(c->device);
  cudaDeviceSynchronize();
  for (void* p : c->opened) cudaIpcCloseMemHandle(p);
  if (c->nccl && c->api) c->api->CommDestroy(c->nccl);
  if (c->xbuf) cudaFree(c->xbuf);
  if (c->h_err) cudaFreeHost(c->h_err);
  for (auto& e : c->ev) if (e) cudaEventDestroy(e);
  delete c;
}

static long long host_hash() {
  char name[256] = {0};
  gethostname(name, sizeof(name) - 1);
  unsigned long long h = 1469598103934665603ull;
  for (const char* p = name; *p; ++p) h = (h ^ (unsigned char)*p) * 1099511628211ull;
  // two containers can share a hostname but not a boot id + pid namespace; good enough to refuse IPC across hosts
  return (long long)(h >> 1);
}

}  // namespace orbfe

using namespace orbfe;

extern "C" {

int orbfe_comm_unique_id(uint8_t* id_out) {
  ORBFE_ARG(id_out);
  NcclApi* api = nccl_api();
  if (!api) return ORBFE_ERR_CUDA;
  NcclUniqueId id;
  ORBFE_NCCL(api, api->GetUniqueId(&id));
  memcpy(id_out, id.internal, ORBFE_COMM_ID_BYTES);
  return ORBFE_OK;
}

int orbfe_comm_create(const uint8_t* id, int nranks, int rank, int device, int max_queries, orbfe_comm** out) {
  ORBFE_ARG(out);
  *out = nullptr;
  ORBFE_ARG(id && nranks >= 1 && nranks <= kMaxRanks && rank >= 0 && rank < nranks && max_queries >= 1);
  int ndev = 0;
  ORBFE_CUDA(cudaGetDeviceCount(&ndev));
  ORBFE_ARG(device >= 0 && device < ndev);
  NcclApi* api = nccl_api();
  if (!api) return ORBFE_ERR_CUDA;
  Comm* c = new (std::nothrow) Comm();
  ORBFE_ARG(c != nullptr);
  c->api = api; c->nranks = nranks; c->rank = rank; c->device = device;
  c->max_queries = (int)align_up((size_t)max_queries, 32);
  DeviceGuard guard(device);
  auto fail = [&](int rc) { comm_free(c); return rc; };
  NcclUniqueId uid;
  static_assert(sizeof(uid.internal) == ORBFE_COMM_ID_BYTES, "id size");
  memcpy(uid.internal, id, ORBFE_COMM_ID_BYTES);
  {
    const int r = api->CommInitRank(&c->nccl, nranks, uid, rank);
    if (r != 0) {
      set_error("comm: ncclCommInitRank(rank %d of %d) -> %d (%s)", rank, nranks, r, api->GetErrorString(r));
      c->nccl = nullptr;
      return fail(ORBFE_ERR_CUDA);
    }
  }
  c->keys_bytes = align_up((size_t)2 * nranks * 2 * c->max_queries * 8, 256);
  const size_t total = c->keys_bytes + align_up((size_t)2 * nranks * 4, 256) + 256;
  if (cudaMalloc(reinterpret_cast<void**>(&c->xbuf), total) != cudaSuccess ||
      cudaMemset(c->xbuf, 0, total) != cudaSuccess ||
      cudaHostAlloc(reinterpret_cast<void**>(&c->h_err), 64, cudaHostAllocMapped) != cudaSuccess) {
    set_error("comm: exchange buffer allocation (%zu bytes) failed: %s", total, cudaGetErrorString(cudaGetLastError()));
    return fail(ORBFE_ERR_CUDA);
  }
  c->h_err[0] = 0;
  for (auto& e : c->ev)
    if (cudaEventCreate(&e) != cudaSuccess) return fail(ORBFE_ERR_CUDA);

  // ---- exchange the cards (one all-gather of nranks x sizeof(PeerCard) bytes) and map the peers' buffers ----
  PeerCard mine;
  memset(&mine, 0, sizeof(mine));
  const bool ipc_ok = cudaIpcGetMemHandle(&mine.xbuf, c->xbuf) == cudaSuccess;
  if (!ipc_ok) cudaGetLastError();
  mine.ptr = reinterpret_cast<unsigned long long>(c->xbuf);
  mine.pid = (long long)getpid();
  mine.host_hash = host_hash();
  mine.device = device;
  mine.pad = ipc_ok ? 1 : 0;
  PeerCard* d_cards = nullptr;
  std::vector<PeerCard> cards(nranks);
  cudaStream_t s = nullptr;
  if (cudaMalloc(reinterpret_cast<void**>(&d_cards), sizeof(PeerCard) * nranks) != cudaSuccess ||
      cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) {
    set_error("comm: set-up allocation failed");
    if (d_cards) cudaFree(d_cards);
    return fail(ORBFE_ERR_CUDA);
  }
  auto gather = [&](void* dev, const void* host_in, void* host_out, size_t each) -> bool {
    if (cudaMemcpyAsync(static_cast<uint8_t*>(dev) + each * rank, host_in, each, cudaMemcpyHostToDevice, s) != cudaSuccess) return false;
    if (api->AllGather(static_cast<uint8_t*>(dev) + each * rank, dev, each, kNcclUint8, c->nccl, s) != 0) return false;
    if (cudaMemcpyAsync(host_out, dev, each * nranks, cudaMemcpyDeviceToHost, s) != cudaSuccess) return false;
    return cudaStreamSynchronize(s) == cudaSuccess;
  };
  bool ok = gather(d_cards, &mine, cards.data(), sizeof(PeerCard));
  int can = ok ? 1 : 0;
  if (ok) {
    const char* force = getenv("ORBFE_COMM_NO_P2P");
    if (force && force[0] == '1') can = 0;
    for (int r = 0; r < nranks && can; ++r) {
      unsigned long long* base = nullptr;
      if (r == rank) {
        base = reinterpret_cast<unsigned long long*>(c->xbuf);
      } else if (cards[r].host_hash != mine.host_hash) {
        can = 0;  // another node: NCCL path
      } else if (cards[r].pid == mine.pid) {  // same process (one thread per GPU): direct pointer + peer access
        int access = 0;
        if (cudaDeviceCanAccessPeer(&access, device, cards[r].device) != cudaSuccess || !access) { can = 0; break; }
        const cudaError_t e = cudaDeviceEnablePeerAccess(cards[r].device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { can = 0; break; }
        cudaGetLastError();
        base = reinterpret_cast<unsigned long long*>(cards[r].ptr);
      } else {
        void* p = nullptr;
        if (!cards[r].pad || cudaIpcOpenMemHandle(&p, cards[r].xbuf, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
          cudaGetLastError();
          can = 0;
          break;
        }
        c->opened.push_back(p);
        base = static_cast<unsigned long long*>(p);
      }
      c->peers.keys[r] = base;
      c->peers.flags[r] = flags_of(reinterpret_cast<uint8_t*>(base), c->keys_bytes);
    }
    // the path is chosen collectively: peer memory only if EVERY rank mapped every peer
    std::vector<int> all(nranks);
    int* d_can = reinterpret_cast<int*>(d_cards);
    ok = gather(d_can, &can, all.data(), sizeof(int));
    if (ok) for (int r = 0; r < nranks; ++r) can &= all[r];
  }
  cudaFree(d_cards);
  cudaStreamDestroy(s);
  if (!ok) {
    set_error("comm: set-up all-gather failed (%s)", cudaGetErrorString(cudaGetLastError()));
    return fail(ORBFE_ERR_CUDA);
  }
  c->p2p = can != 0 && nranks > 1;
  if (!c->p2p) {  // only the own buffer is used
    for (void* p : c->opened) cudaIpcCloseMemHandle(p);
    c->opened.clear();
    c->peers.keys[rank] = reinterpret_cast<unsigned long long*>(c->xbuf);
    c->peers.flags[rank] = flags_of(c->xbuf, c->keys_bytes);
  }
  *out = reinterpret_cast<orbfe_comm*>(c);
  return ORBFE_OK;
}

void orbfe_comm_destroy(orbfe_comm* h) { comm_free(reinterpret_cast<Comm*>(h)); }

int orbfe_comm_info(orbfe_comm* h, int* nranks, int* rank, int* peer_memory, int* nccl_version) {
  ORBFE_ARG(h);
  Comm* c = reinterpret_cast<Comm*>(h);
  if (nranks) *nranks = c->nranks;
  if (rank) *rank = c->rank;
  if (peer_memory) *peer_memory = c->p2p ? 1 : 0;
  if (nccl_version) { *nccl_version = 0; if (c->api->GetVersion) c->api->GetVersion(nccl_version); }
  return ORBFE_OK;
}

int orbfe_comm_status(orbfe_comm* h) {
  ORBFE_ARG(h);
  Comm* c = reinterpret_cast<Comm*>(h);
  if (*reinterpret_cast<volatile unsigned int*>(c->h_err) != 0) {
    set_error("comm: rank %d timed out waiting for a peer's keys (a rank did not make the matching call)", c->rank);
    return ORBFE_ERR_CUDA;
  }
  return ORBFE_OK;
}

int orbfe_comm_set_timing(orbfe_comm* h, int enable) {
  ORBFE_ARG(h);
  reinterpret_cast<Comm*>(h)->timing = enable != 0;
  return ORBFE_OK;
}

int orbfe_comm_last_times(orbfe_comm* h, float* kernel_ms, float* exchange_merge_ms) {
  ORBFE_ARG(h);
  Comm* c = reinterpret_cast<Comm*>(h);
  DeviceGuard guard(c->device);
  ORBFE_CUDA(cudaEventSynchronize(c->ev[2]));
  if (kernel_ms) ORBFE_CUDA(cudaEventElapsedTime(kernel_ms, c->ev[0], c->ev[1]));
  if (exchange_merge_ms) ORBFE_CUDA(cudaEventElapsedTime(exchange_merge_ms, c->ev[1], c->ev[2]));
  return ORBFE_OK;
}

int orbfe_hamming_bruteforce_sharded(orbfe_matcher* hm, orbfe_comm* hc, const uint8_t* d_query, int nq,
                                     const uint8_t* d_map_shard, int nm_local, int64_t index_base, uint64_t* d_best,
                                     uint64_t* d_second, int exchange) {
  ORBFE_ARG(hm && hc && d_query && d_best && d_second && nq >= 1 && nm_local >= 0);
  ORBFE_ARG(nm_local == 0 || d_map_shard);
  ORBFE_ARG(index_base >= 0 && index_base + (int64_t)nm_local <= ((int64_t)1 << 32));
  const bool tensor = (exchange & ORBFE_KERNEL_TENSOR) != 0;  // the shard's top-2 by the tcgen05 kernel
  exchange &= ~ORBFE_KERNEL_TENSOR;
  ORBFE_ARG(exchange == ORBFE_EXCHANGE_AUTO || exchange == ORBFE_EXCHANGE_NCCL || exchange == ORBFE_EXCHANGE_PEER);
  Matcher* m = reinterpret_cast<Matcher*>(hm);
  Comm* c = reinterpret_cast<Comm*>(hc);
  ORBFE_ARG(m->device == c->device);
  if (nq > c->max_queries) {
    set_error("orbfe_hamming_bruteforce_sharded: nq = %d exceeds the communicator's max_queries = %d", nq, c->max_queries);
    return ORBFE_ERR_CAPACITY;
  }
  const bool peer = exchange == ORBFE_EXCHANGE_PEER || (exchange == ORBFE_EXCHANGE_AUTO && c->p2p);
  if (peer && !c->p2p && c->nranks > 1) {
    set_error("orbfe_hamming_bruteforce_sharded: peer-memory exchange asked for, but the ranks cannot map each other");
    return ORBFE_ERR_ARG;
  }
  ORBFE_MATCHER_ENTER(m);
  int slice_len = 0;
  const int slices = nm_local > 0 ? (tensor ? 1 : bruteforce_slices(nq, nm_local, &slice_len)) : 0;
  const int nparts = std::max(slices, 1);
  int rc = ensure_cand(m, (size_t)nparts * nq * 4 + 64);
  if (rc != ORBFE_OK) return rc;
  unsigned long long* pb = reinterpret_cast<unsigned long long*>(m->d_cand);
  unsigned long long* ps = pb + (size_t)nparts * nq;
  cudaStream_t s = m->stream;
  if (c->timing) ORBFE_CUDA(cudaEventRecord(c->ev[0], s));
  if (slices == 0) {  // an empty shard contributes "none"
    ORBFE_CUDA(cudaMemsetAsync(pb, 0xFF, (size_t)nq * 8, s));
    ORBFE_CUDA(cudaMemsetAsync(ps, 0xFF, (size_t)nq * 8, s));
  } else if (tensor) {  // the shard's (best, second) land where one slice's partials would
    rc = matcher_tc_launch(m, d_query, nq, d_map_shard, nm_local, index_base, reinterpret_cast<uint64_t*>(pb),
                           reinterpret_cast<uint64_t*>(ps));
    if (rc != ORBFE_OK) return rc;
  } else {
    launch_hamming_bruteforce(d_query, nq, d_map_shard, nm_local, slice_len, slices, index_base, pb, ps, s);
    m->launches += 1;
  }
  if (c->timing) ORBFE_CUDA(cudaEventRecord(c->ev[1], s));
  const unsigned int epoch = ++c->epoch;
  const int parity = (int)(epoch & 1u);
  unsigned int* my_flags = flags_of(c->xbuf, c->keys_bytes);
  unsigned int* done_counter = my_flags + 2 * c->nranks + 8;
  unsigned int* d_err = nullptr;
  ORBFE_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&d_err), c->h_err, 0));
  const int grid = (nq + 255) / 256;
  PeerTable t = c->peers;
  const bool use_peer = peer && c->nranks > 1;
  if (!use_peer) {  // merge into the own buffer only (the table's other entries are not touched)
    for (int r = 0; r < c->nranks; ++r) { t.keys[r] = c->peers.keys[c->rank]; t.flags[r] = c->peers.flags[c->rank]; }
  }
  merge_push_kernel<<<grid, 256, 0, s>>>(pb, ps, nparts, nq, t, use_peer ? c->nranks : 1, use_peer ? c->rank : 0,
                                         keys_off(c, parity, c->rank, 0), keys_off(c, parity, c->rank, 1),
                                         parity * c->nranks + c->rank, epoch, done_counter);
  if (!use_peer && c->nranks > 1) {
    // in-place all-gather of the (best, second) blocks: block r of this parity belongs to rank r
    unsigned long long* base = reinterpret_cast<unsigned long long*>(c->xbuf) + keys_off(c, parity, 0, 0);
    ORBFE_NCCL(c->api, c->api->AllGather(base + (size_t)c->rank * 2 * c->max_queries, base,
                                         (size_t)2 * c->max_queries, kNcclUint64, c->nccl, s));
  }
  wait_merge_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const unsigned long long*>(c->xbuf),
                                         keys_off(c, parity, 0, 0), c->max_queries, c->nranks, nq,
                                         my_flags + parity * c->nranks, use_peer ? c->nranks : 0, epoch, d_err, wait_timeout_ns(),
                                         reinterpret_cast<unsigned long long*>(d_best),
                                         reinterpret_cast<unsigned long long*>(d_second));
  m->launches += 2;
  if (c->timing) ORBFE_CUDA(cudaEventRecord(c->ev[2], s));
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

}  // extern "C"
