// bz_comm.cu -- the image gather of the row-sharded render behind the C ABI (north_star subsystem 4):
// bz_comm_* wrap an NCCL communicator, bz_gather_rows sends every rank's stripe band to the root over NVLink and assembles the
// frame there.  Rays are independent, so this is the only exchange of the multi-GPU path; a C++ caller of the reference's API
// can shard a frame with nothing but this library and a way to ship 128 bytes to its other ranks.
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
#include "bz_capi.h"
#include "../host/bz_error.h"

cudaError_t bz_launch_assemble(const uint8_t *bands, uint32_t band_rows, uint32_t width, uint32_t height, uint32_t stripe_rows,
                               uint32_t ranks, uint8_t *image, cudaStream_t stream);

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
    api.getErrorString = (decltype(api.getErrorString))load("ncclGetErrorString");
  });
  return api;
}

}

struct bz_comm {
  ncclComm_t comm;
  uint32_t   rank, ranks;
  int        device;
  uint8_t   *d_bands;          // root: the ranks' bands side by side (ranks x band_rows x width x 4), sized on demand
  size_t     bands_bytes;
};

#define BZ_CUDA(call) do { cudaError_t e_ = (call); if(e_ != cudaSuccess) { \
  return bz_fail(BZ_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } } while(0)
#define BZ_NCCL(call) do { ncclResult_t r_ = (call); if(r_ != ncclSuccess) { \
  return bz_fail(BZ_ERR_COMM, std::string(#call) + ": " + nccl().getErrorString(r_)); } } while(0)

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
  *out_comm = out;
  return BZ_OK;
}

int bz_comm_destroy(bz_comm *comm) {
  if(comm == nullptr) { return BZ_OK; }
  cudaSetDevice(comm->device);
  cudaFree(comm->d_bands);
  if(comm->comm != nullptr) { nccl().commDestroy(comm->comm); }
  delete comm;
  return BZ_OK;
}

int bz_gather_rows(bz_comm *comm, const uint8_t *d_band, uint32_t width, uint32_t height, uint32_t stripe_rows, uint32_t root,
                   uint8_t *d_rgba, void *stream) {
  if(comm == nullptr) { return bz_fail(BZ_ERR_INVALID, "comm is null"); }
  if(stripe_rows == 0u || root >= comm->ranks) { return bz_fail(BZ_ERR_INVALID, "bad stripe layout or root"); }
  cudaStream_t const st = (cudaStream_t)stream;
  size_t const row_bytes = (size_t)width * 4u;
  uint32_t const own_rows = bz_striped_rows(height, stripe_rows, comm->rank, comm->ranks);
  if(own_rows != 0u && d_band == nullptr) { return bz_fail(BZ_ERR_INVALID, "d_band is null"); }
  BZ_CUDA(cudaSetDevice(comm->device));
  if(comm->rank != root) {
    if(own_rows != 0u) { BZ_NCCL(nccl().send(d_band, own_rows * row_bytes, ncclUint8, (int)root, comm->comm, st)); }
    return BZ_OK;
  }
  if((uint64_t)width * height != 0u && d_rgba == nullptr) { return bz_fail(BZ_ERR_INVALID, "d_rgba is null on the root"); }
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

}  // extern "C"
