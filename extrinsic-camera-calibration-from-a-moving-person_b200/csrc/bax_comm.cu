// All-reduces of the packed reduce buffers INSIDE the library (bax_allreduce_attach): a frame-sharded
// run sums a handful of small buffers over the ranks between the kernel phases of every linearisation and damped solve.
// Round 1 returned to the host language for each of them (torch.distributed between ctypes calls), which kept the
// sharded solver out of the one-call-per-iteration native driver and cost ~0.25 ms of interpreter time and rank skew
// per iteration at 8 GPUs.  With a communicator attached, bax_trf_begin / _iterate / _solve issue the collectives
// themselves, on the compute stream, between the phases.
//
// NCCL is bound at run time (dlopen of the libnccl.so.2 already in the process — PyTorch's — else the system's), so
// libbax.so has no link-time dependency on it and single-GPU users never touch it.
#include "bax_internal.cuh"
#include <dlfcn.h>
#include <string.h>

namespace {

typedef struct { char internal[128]; } NcclUniqueId;            // ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES = 128)
typedef void* NcclComm;
enum { kNcclSum = 0, kNcclMax = 2, kNcclFloat64 = 8 };          // ncclRedOp_t / ncclDataType_t values

struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

NcclApi* api() {
  static NcclApi a;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW);
    if (lib) {
      a.GetUniqueId = (int (*)(NcclUniqueId*))dlsym(lib, "ncclGetUniqueId");
      a.CommInitRank = (int (*)(NcclComm*, int, NcclUniqueId, int))dlsym(lib, "ncclCommInitRank");
      a.CommDestroy = (int (*)(NcclComm))dlsym(lib, "ncclCommDestroy");
      a.AllReduce = (int (*)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t))dlsym(lib, "ncclAllReduce");
      a.GroupStart = (int (*)())dlsym(lib, "ncclGroupStart");
      a.GroupEnd = (int (*)())dlsym(lib, "ncclGroupEnd");
      a.GetErrorString = (const char* (*)(int))dlsym(lib, "ncclGetErrorString");
      if (a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllReduce && a.GroupStart && a.GroupEnd) a.lib = lib;
    }
  }
  return a.lib ? &a : nullptr;
}

int nccl_fail(const char* what, int rc) {
  NcclApi* a = api();
  bax_set_error("%s: NCCL error %d (%s)", what, rc, a && a->GetErrorString ? a->GetErrorString(rc) : "?");
  return BAX_E_CUDA;
}

}  // namespace

extern "C" int bax_comm_unique_id(void* id128_host) {
  NcclApi* a = api();
  if (!a) { bax_set_error("bax_comm_unique_id: libnccl.so.2 not found"); return BAX_E_STATE; }
  if (!id128_host) { bax_set_error("bax_comm_unique_id: bad argument"); return BAX_E_ARG; }
  NcclUniqueId id;
  int rc = a->GetUniqueId(&id);
  if (rc) return nccl_fail("ncclGetUniqueId", rc);
  memcpy(id128_host, &id, sizeof(id));
  return BAX_OK;
}

extern "C" int bax_comm_create(int device, const void* id128_host, int rank, int world, void** comm_out) {
  NcclApi* a = api();
  if (!a) { bax_set_error("bax_comm_create: libnccl.so.2 not found"); return BAX_E_STATE; }
  if (!id128_host || !comm_out || world < 1 || rank < 0 || rank >= world) { bax_set_error("bax_comm_create: bad argument"); return BAX_E_ARG; }
  BAX_CUDA(cudaSetDevice(device));
  NcclUniqueId id;
  memcpy(&id, id128_host, sizeof(id));
  NcclComm c = nullptr;
  int rc = a->CommInitRank(&c, world, id, rank);
  if (rc) return nccl_fail("ncclCommInitRank", rc);
  *comm_out = c;
  return BAX_OK;
}

extern "C" int bax_comm_destroy(void* comm) {
  NcclApi* a = api();
  if (a && comm) a->CommDestroy((NcclComm)comm);
  return BAX_OK;
}

extern "C" int bax_allreduce_attach(bax_handle* h, void* nccl_comm) {
  if (!h) { bax_set_error("bax_allreduce_attach: bad argument"); return BAX_E_ARG; }
  if (nccl_comm && !api()) { bax_set_error("bax_allreduce_attach: libnccl.so.2 not found"); return BAX_E_STATE; }
  h->comm = nccl_comm;                                        // borrowed; NULL detaches
  return BAX_OK;
}

// In-place sum (or max) of one reduce buffer over the ranks, on stream s.  No-op without a communicator.
int bax_allreduce(bax_handle* h, int which, bool op_max, cudaStream_t s) {
  if (!h->comm) return BAX_OK;
  NcclApi* a = api();
  double* p = h->buf[which == BAX_BUF_RHS2_HEAD ? BAX_BUF_RHS2 : which];
  const int64_t n = which == BAX_BUF_RHS2_HEAD ? 6 * (int64_t)h->C + h->B : h->buf_count[which];
  int rc = a->AllReduce(p, p, (size_t)n, kNcclFloat64, op_max ? kNcclMax : kNcclSum, (NcclComm)h->comm, s);
  if (rc) return nccl_fail("ncclAllReduce", rc);
  h->n_collectives += 1;
  return BAX_OK;
}

// Two buffers in ONE group call (one launch): `which_sum` summed, `which_max` max-reduced.
int bax_allreduce_pair(bax_handle* h, int which_sum, int which_max, cudaStream_t s) {
  if (!h->comm) return BAX_OK;
  NcclApi* a = api();
  int rc = a->GroupStart();
  if (rc) return nccl_fail("ncclGroupStart", rc);
  rc = a->AllReduce(h->buf[which_sum], h->buf[which_sum], (size_t)h->buf_count[which_sum], kNcclFloat64, kNcclSum, (NcclComm)h->comm, s);
  if (!rc) rc = a->AllReduce(h->buf[which_max], h->buf[which_max], (size_t)h->buf_count[which_max], kNcclFloat64, kNcclMax, (NcclComm)h->comm, s);
  int rc2 = a->GroupEnd();
  if (rc || rc2) return nccl_fail("ncclAllReduce (group)", rc ? rc : rc2);
  h->n_collectives += 1;
  return BAX_OK;
}

extern "C" int bax_comm_stats(bax_handle* h, int64_t* out_host) {
  if (!h || !out_host) { bax_set_error("bax_comm_stats: bad argument"); return BAX_E_ARG; }
  out_host[0] = h->comm ? 1 : 0; out_host[1] = h->n_collectives;
  return BAX_OK;
}
