// bz_comm.cu -- the image gather of the row-sharded render behind the C ABI (north_star subsystem 4):
// bz_comm_* wrap an NCCL communicator, bz_gather_rows brings every rank's stripe band to the root over NVLink and leaves the
// assembled frame there.  Rays are independent, so this is the only exchange of the multi-GPU path; a C++ caller of the reference's
// API can shard a frame with nothing but this library and a way to ship 128 bytes to its other ranks.
//
// Two transports, same result:
//   BZ_GATHER_PEER (default)  the root owns an image buffer that every other rank maps into its own process (CUDA IPC); a rank's
//                             bz_scatter_band_kernel STORES its rows straight into their place in the root's buffer over NVLink --
//                             transfer and assembly are one kernel per rank, all ranks in parallel.  "My rows of call k are there"
//                             and "the frame of call k has been taken out of the buffer" are generation counters in the root's
//                             buffer, written and polled across NVLink by the same kernels: no NCCL call per frame (NCCL carries
//                             the IPC handle once per frame size)
//   BZ_GATHER_NCCL            every band travels by ncclSend / ncclRecv into a staging buffer on the root, a second pass puts the
//                             rows in place (round 2's first form; kept as the fallback when IPC mapping is not available)
//
// NCCL is loaded at run time (dlopen "libnccl.so.2": the copy already in the process when PyTorch is, the system's otherwise),
// so the library itself loads and runs on machines without it; only bz_comm_* need it and say so when it is missing.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>
#include "bz_capi.h"
#include "../host/bz_error.h"

cudaError_t bz_launch_assemble(const uint8_t *bands, uint32_t band_rows, uint32_t width, uint32_t height, uint32_t stripe_rows,
                               uint32_t ranks, uint8_t *image, cudaStream_t stream);
cudaError_t bz_launch_scatter_band(const uint8_t *band, uint32_t rows, uint32_t width, uint32_t stripe_rows, uint32_t rank, uint32_t ranks,
                                   uint8_t *image, cudaStream_t stream);
cudaError_t bz_launch_scatter_band_flagged(const uint8_t *band, uint32_t rows, uint32_t width, uint32_t stripe_rows, uint32_t rank, uint32_t ranks, uint8_t *image,
                                           const uint32_t *free_word, uint32_t need_free, uint32_t *done_word, uint32_t done_value, uint32_t *finished, cudaStream_t stream);
cudaError_t bz_launch_wait_ranks(const uint32_t *done_words, uint32_t ranks, uint32_t root, uint32_t value, cudaStream_t stream);
cudaError_t bz_launch_publish_word(uint32_t *word, uint32_t value, cudaStream_t stream);

namespace {

struct NcclApi {
  void *handle = nullptr;
  decltype(&ncclGetUniqueId) getUniqueId = nullptr;
  decltype(&ncclCommInitRank) commInitRank = nullptr;
  decltype(&ncclCommDestroy) commDestroy = nullptr;
  decltype(&ncclSend) send = nullptr;
  decltype(&ncclRecv) recv = nullptr;
  decltype(&ncclGroupStart) groupStart = nullptr;
  decltype(&ncclGroupEnd) groupEnd = nullptr;
  decltype(&ncclBroadcast) broadcast = nullptr;
  decltype(&ncclAllReduce) allReduce = nullptr;
  decltype(&ncclGetErrorString) getErrorString = nullptr;
  std::string problem;
};

NcclApi &nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    for(const char *name : { "libnccl.so.2", "libnccl.so" }) {
      api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if(api.handle != nullptr) { break; }
    }
    if(api.handle == nullptr) { api.problem = "NCCL is not available (dlopen libnccl.so.2 failed)"; return; }
    auto load = [&](const char *symbol) { void *p = dlsym(api.handle, symbol); if(p == nullptr && api.problem.empty()) { api.problem = std::string("NCCL lacks ") + symbol; } return p; };
    api.getUniqueId = (decltype(api.getUniqueId))load("ncclGetUniqueId");
    api.commInitRank = (decltype(api.commInitRank))load("ncclCommInitRank");
    api.commDestroy = (decltype(api.commDestroy))load("ncclCommDestroy");
    api.send = (decltype(api.send))load("ncclSend");
    api.recv = (decltype(api.recv))load("ncclRecv");
    api.groupStart = (decltype(api.groupStart))load("ncclGroupStart");
    api.groupEnd = (decltype(api.groupEnd))load("ncclGroupEnd");
    api.broadcast = (decltype(api.broadcast))load("ncclBroadcast");
    api.allReduce = (decltype(api.allReduce))load("ncclAllReduce");
    api.getErrorString = (decltype(api.getErrorString))load("ncclGetErrorString");
  });
  return api;
}

}

struct bz_comm {
  ncclComm_t comm;
  uint32_t   rank, ranks;
  int        device;
  uint32_t   mode;             // BZ_GATHER_PEER / BZ_GATHER_NCCL as asked for
  bool       peer_failed;      // the ranks agreed that IPC mapping does not work here: NCCL transport from then on
  // BZ_GATHER_NCCL
  uint8_t   *d_bands;          // root: the ranks' bands side by side (ranks x band_rows x width x 4), sized on demand
  size_t     bands_bytes;
  // BZ_GATHER_PEER
  uint8_t   *d_image;          // the root's image buffer: its own allocation on the root, the IPC mapping of it elsewhere
  size_t     image_bytes;      // size the current buffer was agreed for (0: none yet)
  uint32_t   image_root;
  uint32_t   calls;            // gather calls since the buffer was set up: the generation the flags count in
  uint32_t  *d_token;          // 64 device words of this rank: [0] outcome of the set-up vote, [3] CTAs of the scatter kernel that are done,
                               // [4 ..] the IPC handle on its way
  std::vector<void*> retired;  // root: earlier image buffers (other ranks may still map them; freed with the communicator)
};

#define BZ_CUDA(call) do { cudaError_t e_ = (call); if(e_ != cudaSuccess) { \
  return bz_fail(BZ_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } } while(0)
#define BZ_NCCL(call) do { ncclResult_t r_ = (call); if(r_ != ncclSuccess) { \
  return bz_fail(BZ_ERR_COMM, std::string(#call) + ": " + nccl().getErrorString(r_)); } } while(0)

// ---- BZ_GATHER_NCCL: bands by ncclSend / ncclRecv into a staging buffer on the root, then one pass that puts the rows in place ----
static int bz_gather_nccl(bz_comm *comm, const uint8_t *d_band, uint32_t own_rows, uint32_t width, uint32_t height, uint32_t stripe_rows,
                          uint32_t root, uint8_t *d_rgba, cudaStream_t st) {
  size_t const row_bytes = (size_t)width * 4u;
  if(comm->rank != root) {
    if(own_rows != 0u) { BZ_NCCL(nccl().send(d_band, own_rows * row_bytes, ncclUint8, (int)root, comm->comm, st)); }
    return BZ_OK;
  }
  uint32_t band_rows = 0u;
  for(uint32_t r = 0u; r < comm->ranks; ++r) { band_rows = std::max(band_rows, bz_striped_rows(height, stripe_rows, r, comm->ranks)); }
  size_t const band_bytes = (size_t)band_rows * row_bytes;
  if(band_bytes * comm->ranks > comm->bands_bytes) {
    if(comm->d_bands != nullptr) { BZ_CUDA(cudaFree(comm->d_bands)); comm->d_bands = nullptr; comm->bands_bytes = 0u; }
    BZ_CUDA(cudaMalloc((void**)&comm->d_bands, band_bytes * comm->ranks));
    comm->bands_bytes = band_bytes * comm->ranks;
  }
  // every other rank's band straight into its slot, the root's own by a device copy, then one pass that puts the rows in place
  BZ_NCCL(nccl().groupStart());
  for(uint32_t r = 0u; r < comm->ranks; ++r) {
    uint32_t const rows = bz_striped_rows(height, stripe_rows, r, comm->ranks);
    if(r == root || rows == 0u) { continue; }
    ncclResult_t const result = nccl().recv(comm->d_bands + (size_t)r * band_bytes, rows * row_bytes, ncclUint8, (int)r, comm->comm, st);
    if(result != ncclSuccess) { nccl().groupEnd(); return bz_fail(BZ_ERR_COMM, std::string("ncclRecv: ") + nccl().getErrorString(result)); }
  }
  BZ_NCCL(nccl().groupEnd());
  if(own_rows != 0u) { BZ_CUDA(cudaMemcpyAsync(comm->d_bands + (size_t)root * band_bytes, d_band, own_rows * row_bytes, cudaMemcpyDeviceToDevice, st)); }
  BZ_CUDA(bz_launch_assemble(comm->d_bands, band_rows, width, height, stripe_rows, comm->ranks, d_rgba, st));
  return BZ_OK;
}

// ---- BZ_GATHER_PEER ----
// Layout of the root's buffer: the frame (rounded up to 256 bytes), then a control block of generation counters:
//   word 0           frames taken out of the buffer so far (written by the root, polled by the others before they store the next frame)
//   word 64 + r      last call whose rows rank r has stored (written by rank r, polled by the root)
static size_t bz_peer_frame_bytes(size_t bytes) { return (bytes + 255u) / 256u * 256u; }
static size_t bz_peer_control_bytes(uint32_t ranks) { return ((size_t)(64u + ranks) * sizeof(uint32_t) + 255u) / 256u * 256u; }

// The root's buffer for frames of `bytes` bytes, mapped on every rank.  Collective and synchronous, but only when the frame size or
// the root changes (every rank sees the same arguments, so all of them take this path together): the root allocates, its IPC handle
// travels by ncclBroadcast, the others open it, and an all-reduce of the outcome decides for everybody whether the peer transport is used.
static int bz_peer_prepare(bz_comm *comm, size_t bytes, uint32_t root, cudaStream_t st) {
  if(comm->image_bytes == bytes && comm->image_root == root) { return BZ_OK; }
  static_assert(sizeof(cudaIpcMemHandle_t) <= 60u * sizeof(uint32_t), "the handle travels in the token words");
  BZ_CUDA(cudaStreamSynchronize(st));
  if(comm->d_image != nullptr) {
    if(comm->rank == comm->image_root) { comm->retired.push_back(comm->d_image); }      // (others may still map it)
    else { BZ_CUDA(cudaIpcCloseMemHandle(comm->d_image)); }
    comm->d_image = nullptr;
  }
  comm->image_bytes = 0u;
  comm->calls = 0u;
  uint32_t ok = (comm->ranks <= 1024u) ? 1u : 0u;             // (bz_wait_ranks_kernel: one thread per rank)
  size_t const total = bz_peer_frame_bytes(bytes) + bz_peer_control_bytes(comm->ranks);
  cudaIpcMemHandle_t handle;
  std::memset(&handle, 0, sizeof(handle));
  if(comm->rank == root && ok != 0u) {
    if(cudaMalloc((void**)&comm->d_image, total) != cudaSuccess) { cudaGetLastError(); comm->d_image = nullptr; ok = 0u; }
    else if(cudaMemset(comm->d_image + bz_peer_frame_bytes(bytes), 0, bz_peer_control_bytes(comm->ranks)) != cudaSuccess) { cudaGetLastError(); ok = 0u; }
    else if(comm->ranks > 1u && cudaIpcGetMemHandle(&handle, comm->d_image) != cudaSuccess) { cudaGetLastError(); ok = 0u; }
  }
  BZ_CUDA(cudaMemsetAsync(comm->d_token, 0, 4u * sizeof(uint32_t), st));
  if(comm->ranks > 1u) {
    if(comm->rank == root) { BZ_CUDA(cudaMemcpyAsync(comm->d_token + 4, &handle, sizeof(handle), cudaMemcpyHostToDevice, st)); }
    BZ_NCCL(nccl().broadcast(comm->d_token + 4, comm->d_token + 4, sizeof(handle), ncclUint8, (int)root, comm->comm, st));
    if(comm->rank != root) {
      BZ_CUDA(cudaMemcpyAsync(&handle, comm->d_token + 4, sizeof(handle), cudaMemcpyDeviceToHost, st));
      BZ_CUDA(cudaStreamSynchronize(st));
      void *mapped = nullptr;
      if(ok != 0u && cudaIpcOpenMemHandle(&mapped, handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0u; }
      else { comm->d_image = (uint8_t*)mapped; }
    }
    // everybody, or nobody
    BZ_CUDA(cudaMemcpyAsync(comm->d_token, &ok, sizeof(ok), cudaMemcpyHostToDevice, st));
    BZ_NCCL(nccl().allReduce(comm->d_token, comm->d_token, 1, ncclUint32, ncclMin, comm->comm, st));
    BZ_CUDA(cudaMemcpyAsync(&ok, comm->d_token, sizeof(ok), cudaMemcpyDeviceToHost, st));
    BZ_CUDA(cudaStreamSynchronize(st));
  }
  if(ok == 0u) {
    if(comm->d_image != nullptr) {
      if(comm->rank == root) { cudaFree(comm->d_image); } else { cudaIpcCloseMemHandle(comm->d_image); }
      cudaGetLastError();
      comm->d_image = nullptr;
    }
    comm->peer_failed = true;
    return BZ_OK;
  }
  comm->image_bytes = bytes;
  comm->image_root = root;
  return BZ_OK;
}

// Call k of the communicator on this buffer.  A rank other than the root: ONE kernel that waits (normally not at all) until the root has
// taken frame k - 1 out of the buffer, stores the band's rows into their place over NVLink and publishes "rows of call k are there".
// The root: its own rows (a local scatter), a one-CTA kernel that waits for every rank's word, the copy of the frame to the caller's
// d_rgba, and "frame k is out of the buffer".  Nobody waits at the end of a call; the flags only ever grow.
static int bz_gather_peer(bz_comm *comm, const uint8_t *d_band, uint32_t own_rows, uint32_t width, uint32_t height, uint32_t stripe_rows,
                          uint32_t root, uint8_t *d_rgba, cudaStream_t st) {
  uint32_t const call = ++comm->calls;
  size_t const bytes = (size_t)width * height * 4u;
  uint32_t *control = (uint32_t*)(comm->d_image + bz_peer_frame_bytes(bytes));
  if(comm->rank != root) {
    BZ_CUDA(bz_launch_scatter_band_flagged(d_band, own_rows, width, stripe_rows, comm->rank, comm->ranks, comm->d_image,
                                           control, call - 1u, control + 64u + comm->rank, call, comm->d_token + 3, st));
    return BZ_OK;
  }
  BZ_CUDA(bz_launch_scatter_band(d_band, own_rows, width, stripe_rows, comm->rank, comm->ranks, comm->d_image, st));
  if(comm->ranks > 1u) { BZ_CUDA(bz_launch_wait_ranks(control + 64u, comm->ranks, root, call, st)); }
  BZ_CUDA(cudaMemcpyAsync(d_rgba, comm->d_image, bytes, cudaMemcpyDeviceToDevice, st));
  if(comm->ranks > 1u) { BZ_CUDA(bz_launch_publish_word(control, call, st)); }
  return BZ_OK;
}

extern "C" {

int bz_comm_unique_id(uint8_t out_id[BZ_COMM_ID_BYTES]) {
  if(out_id == nullptr) { return bz_fail(BZ_ERR_INVALID, "out_id is null"); }
  static_assert(BZ_COMM_ID_BYTES == NCCL_UNIQUE_ID_BYTES, "the id is NCCL's");
  if(!nccl().problem.empty()) { return bz_fail(BZ_ERR_COMM, nccl().problem); }
  ncclUniqueId id;
  BZ_NCCL(nccl().getUniqueId(&id));
  std::memcpy(out_id, id.internal, BZ_COMM_ID_BYTES);
  return BZ_OK;
}

int bz_comm_create(const uint8_t id[BZ_COMM_ID_BYTES], uint32_t rank, uint32_t ranks, int device, bz_comm **out_comm) {
  if(out_comm == nullptr) { return bz_fail(BZ_ERR_INVALID, "out_comm is null"); }
  *out_comm = nullptr;
  if(id == nullptr || ranks == 0u || rank >= ranks) { return bz_fail(BZ_ERR_INVALID, "bad id or rank"); }
  if(!nccl().problem.empty()) { return bz_fail(BZ_ERR_COMM, nccl().problem); }
  int count = 0;
  if(cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { cudaGetLastError(); return bz_fail(BZ_ERR_NO_DEVICE, "no usable CUDA device (there is no CPU path)"); }
  if(device < 0 || device >= count) { return bz_fail(BZ_ERR_INVALID, "device index out of range"); }
  BZ_CUDA(cudaSetDevice(device));
  ncclUniqueId unique;
  std::memcpy(unique.internal, id, BZ_COMM_ID_BYTES);
  ncclComm_t comm = nullptr;
  BZ_NCCL(nccl().commInitRank(&comm, (int)ranks, unique, (int)rank));      // collective: every rank of the id calls it
  bz_comm *out = new bz_comm();
  out->comm = comm;
  out->rank = rank;
  out->ranks = ranks;
  out->device = device;
  out->mode = BZ_GATHER_PEER;
  if(cudaMalloc((void**)&out->d_token, 64u * sizeof(uint32_t)) != cudaSuccess || cudaMemset(out->d_token, 0, 64u * sizeof(uint32_t)) != cudaSuccess) {
    cudaGetLastError();
    nccl().commDestroy(comm);
    delete out;
    return bz_fail(BZ_ERR_CUDA, "cudaMalloc of the communicator's token words failed");
  }
  *out_comm = out;
  return BZ_OK;
}

int bz_comm_set_gather(bz_comm *comm, uint32_t mode) {
  if(comm == nullptr) { return bz_fail(BZ_ERR_INVALID, "comm is null"); }
  if(mode != BZ_GATHER_PEER && mode != BZ_GATHER_NCCL) { return bz_fail(BZ_ERR_INVALID, "unknown gather mode"); }
  comm->mode = mode;           // (every rank of the communicator must ask for the same mode)
  return BZ_OK;
}

int bz_comm_gather_mode(const bz_comm *comm, uint32_t *out_mode) {
  if(comm == nullptr || out_mode == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  *out_mode = (comm->mode == BZ_GATHER_PEER && !comm->peer_failed) ? BZ_GATHER_PEER : BZ_GATHER_NCCL;
  return BZ_OK;
}

int bz_comm_destroy(bz_comm *comm) {
  if(comm == nullptr) { return BZ_OK; }
  cudaSetDevice(comm->device);
  cudaDeviceSynchronize();
  cudaFree(comm->d_bands);
  cudaFree(comm->d_token);
  // the ranks that map the root's frame buffer let go of it BEFORE the communicator's teardown (where NCCL makes the ranks meet), the
  // root frees it after
  bool const owner = (comm->d_image != nullptr && comm->rank == comm->image_root);
  if(comm->d_image != nullptr && !owner) { cudaIpcCloseMemHandle(comm->d_image); }
  cudaGetLastError();
  if(comm->comm != nullptr) { nccl().commDestroy(comm->comm); }
  if(owner) { cudaFree(comm->d_image); }
  for(void *p : comm->retired) { cudaFree(p); }
  cudaGetLastError();
  delete comm;
  return BZ_OK;
}

int bz_gather_rows(bz_comm *comm, const uint8_t *d_band, uint32_t width, uint32_t height, uint32_t stripe_rows, uint32_t root,
                   uint8_t *d_rgba, void *stream) {
  if(comm == nullptr) { return bz_fail(BZ_ERR_INVALID, "comm is null"); }
  if(stripe_rows == 0u || root >= comm->ranks) { return bz_fail(BZ_ERR_INVALID, "bad stripe layout or root"); }
  cudaStream_t const st = (cudaStream_t)stream;
  uint32_t const own_rows = bz_striped_rows(height, stripe_rows, comm->rank, comm->ranks);
  if(own_rows != 0u && d_band == nullptr) { return bz_fail(BZ_ERR_INVALID, "d_band is null"); }
  if(comm->rank == root && (uint64_t)width * height != 0u && d_rgba == nullptr) { return bz_fail(BZ_ERR_INVALID, "d_rgba is null on the root"); }
  BZ_CUDA(cudaSetDevice(comm->device));
  if((uint64_t)width * height == 0u) { return BZ_OK; }
  if(comm->mode == BZ_GATHER_PEER && !comm->peer_failed) {
    int rc = bz_peer_prepare(comm, (size_t)width * height * 4u, root, st);
    if(rc != BZ_OK) { return rc; }
  }
  if(comm->mode == BZ_GATHER_PEER && !comm->peer_failed) { return bz_gather_peer(comm, d_band, own_rows, width, height, stripe_rows, root, d_rgba, st); }
  return bz_gather_nccl(comm, d_band, own_rows, width, height, stripe_rows, root, d_rgba, st);
}

}  // extern "C"
