/*
 * Minimal run-time binding to NCCL (libnccl.so.2) through dlopen, so that the single-GPU path has no
 * link-time dependency on NCCL and so that, inside a Python process that already imported torch, the
 * NCCL that torch loaded is the one used (same SONAME -> same mapping).
 * Only the point-to-point subset the z-slab halo exchange needs is bound.
 */
#ifndef FDTD_NCCL_DL_H
#define FDTD_NCCL_DL_H

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stddef.h>

namespace fdtd
{

struct NcclUniqueId { char internal[128]; };
typedef void *NcclComm;

struct NcclApi
{
  void *lib;
  int (*GetUniqueId) (NcclUniqueId *);
  int (*CommInitRank) (NcclComm *, int, NcclUniqueId, int);
  int (*CommDestroy) (NcclComm);
  int (*Send) (const void *, size_t, int, int, NcclComm, cudaStream_t);
  int (*Recv) (void *, size_t, int, int, NcclComm, cudaStream_t);
  int (*GroupStart) ();
  int (*GroupEnd) ();
  const char *(*GetErrorString) (int);
};

enum { NCCL_INT8 = 0 }; /* ncclInt8 / ncclChar */

struct NcclLoad
{
  NcclApi api;
  const char *err;
  bool ok;
  NcclLoad () : err (""), ok (false)
  {
    api.lib = dlopen ("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!api.lib) api.lib = dlopen ("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!api.lib) { err = "cannot dlopen libnccl.so.2"; return; }
#define FDTD_NCCL_SYM(field, name) \
  *(void **) (&api.field) = dlsym (api.lib, name); \
  if (!api.field) err = "libnccl lacks " name;
    FDTD_NCCL_SYM (GetUniqueId, "ncclGetUniqueId")
    FDTD_NCCL_SYM (CommInitRank, "ncclCommInitRank")
    FDTD_NCCL_SYM (CommDestroy, "ncclCommDestroy")
    FDTD_NCCL_SYM (Send, "ncclSend")
    FDTD_NCCL_SYM (Recv, "ncclRecv")
    FDTD_NCCL_SYM (GroupStart, "ncclGroupStart")
    FDTD_NCCL_SYM (GroupEnd, "ncclGroupEnd")
    FDTD_NCCL_SYM (GetErrorString, "ncclGetErrorString")
#undef FDTD_NCCL_SYM
    ok = err[0] == 0;
  }
};

/* Returns NULL (and sets *why) if NCCL cannot be loaded.  The function-local static is initialised exactly once, also
 * when several threads create solvers at the same time (C++11 guarantees it). */
inline const NcclApi *nccl_api (const char **why)
{
  static const NcclLoad load;
  if (why) *why = load.err;
  return load.ok ? &load.api : NULL;
}

} /* namespace fdtd */

#endif /* FDTD_NCCL_DL_H */
