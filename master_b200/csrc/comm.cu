// comm.cu — the one collective of the path behind the boundary: b2pt_comm_* / b2pt_reduce_accum, and the device-resident film.
//
// The reference has no distributed mode; its multi-machine workflow renders independently and adds the EXR accumulators offline:
// merge_exr (Options.cpp:1340-1409) computes out = a + b on all four channels, the fourth being the sample count.  Here every GPU
// renders its own sample indices of a replicated scene into a device-resident dvec4 film and the films are added with ONE
// ncclAllReduce / ncclReduce (sum, float64) over NVLink — exactly that arithmetic.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy a host process has already loaded, e.g. torch's, or the system's), so
// libb2pt.so itself has no link-time dependency on it and single-GPU users never touch it.  Only the six entry points below are
// used; their prototypes are restated from nccl.h (stable since NCCL 2.0).
#include <dlfcn.h>
#include <string.h>

#include <string>

#include "render_common.cuh"

namespace b2pt {
namespace {

struct NcclUniqueId { char internal[128]; };  // ncclUniqueId
typedef void* NcclComm;
enum { kNcclSuccess = 0, kNcclFloat64 = 8, kNcclSum = 0 };

struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*Reduce)(const void*, void*, size_t, int, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  std::string error;
};

NcclApi* nccl() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return &api;
  tried = true;
  const char* env = getenv("B2PT_NCCL_LIB");
  const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    if (!n || !*n) continue;
    api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.handle) break;
  }
  if (!api.handle) { api.error = std::string("cannot load libnccl.so.2 (set B2PT_NCCL_LIB): ") + (dlerror() ? dlerror() : ""); return &api; }
  auto sym = [&](const char* s) { void* p = dlsym(api.handle, s); if (!p) api.error = std::string("libnccl: missing symbol ") + s; return p; };
  api.GetUniqueId = (int (*)(NcclUniqueId*))sym("ncclGetUniqueId");
  api.CommInitRank = (int (*)(NcclComm*, int, NcclUniqueId, int))sym("ncclCommInitRank");
  api.AllReduce = (int (*)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t))sym("ncclAllReduce");
  api.Reduce = (int (*)(const void*, void*, size_t, int, int, int, NcclComm, cudaStream_t))sym("ncclReduce");
  api.CommDestroy = (int (*)(NcclComm))sym("ncclCommDestroy");
  api.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
  return &api;
}

int nccl_fail(NcclApi* a, int rc, const char* what) {
  set_error(std::string(what) + " failed: " + (a->GetErrorString ? a->GetErrorString(rc) : "NCCL error"));
  return B2PT_ERR_CUDA;
}

}  // namespace
}  // namespace b2pt

struct b2pt_comm_s {
  b2pt::NcclComm comm = nullptr;
  int nranks = 1, rank = 0, device = 0;
};

struct b2pt_film_s {
  double* data = nullptr;
  uint32_t width = 0, height = 0;
  int device = 0;
};

extern "C" {

int b2pt_comm_unique_id(void* id) {
  using namespace b2pt;
  B2_CHECK(id, B2PT_ERR_INVALID, "b2pt_comm_unique_id: null argument");
  NcclApi* a = nccl();
  B2_CHECK(a->error.empty(), B2PT_ERR_CUDA, a->error);
  NcclUniqueId u;
  const int rc = a->GetUniqueId(&u);
  if (rc != kNcclSuccess) return nccl_fail(a, rc, "ncclGetUniqueId");
  memcpy(id, &u, sizeof u);
  return B2PT_OK;
}

int b2pt_comm_init(int nranks, int rank, const void* id, b2pt_comm_t* out) {
  using namespace b2pt;
  B2_CHECK(out && nranks >= 1 && rank >= 0 && rank < nranks, B2PT_ERR_INVALID, "b2pt_comm_init: bad rank / nranks");
  b2pt_comm_s* c = new b2pt_comm_s();
  c->nranks = nranks; c->rank = rank;
  if (cudaGetDevice(&c->device) != cudaSuccess) { cudaGetLastError(); delete c; B2_CHECK(false, B2PT_ERR_NO_DEVICE, "b2pt_comm_init: no CUDA device"); }
  if (nranks > 1) {  // one rank needs no communicator: the reduce is the identity
    B2_CHECK(id, B2PT_ERR_INVALID, "b2pt_comm_init: null id");
    NcclApi* a = nccl();
    if (!a->error.empty()) { delete c; B2_CHECK(false, B2PT_ERR_CUDA, a->error); }
    NcclUniqueId u;
    memcpy(&u, id, sizeof u);
    const int rc = a->CommInitRank(&c->comm, nranks, u, rank);
    if (rc != kNcclSuccess) { delete c; return nccl_fail(a, rc, "ncclCommInitRank"); }
  }
  *out = c;
  return B2PT_OK;
}

int b2pt_reduce_accum(b2pt_comm_t comm, double* accum_dev, size_t num_doubles, int root, void* stream) {
  using namespace b2pt;
  B2_CHECK(comm && (accum_dev || num_doubles == 0), B2PT_ERR_INVALID, "b2pt_reduce_accum: null argument");
  B2_CHECK(root >= -1 && root < comm->nranks, B2PT_ERR_INVALID, "b2pt_reduce_accum: bad root");
  if (comm->nranks == 1 || num_doubles == 0) return B2PT_OK;
  NcclApi* a = nccl();
  const int rc = root < 0 ? a->AllReduce(accum_dev, accum_dev, num_doubles, kNcclFloat64, kNcclSum, comm->comm, (cudaStream_t)stream)
                          : a->Reduce(accum_dev, accum_dev, num_doubles, kNcclFloat64, kNcclSum, root, comm->comm, (cudaStream_t)stream);
  if (rc != kNcclSuccess) return nccl_fail(a, rc, root < 0 ? "ncclAllReduce" : "ncclReduce");
  return B2PT_OK;
}

void b2pt_comm_destroy(b2pt_comm_t comm) {
  if (!comm) return;
  if (comm->comm) b2pt::nccl()->CommDestroy(comm->comm);
  delete comm;
}

// ---- the device-resident film: subimage_view_t's dvec4 buffer (ImageView.hpp:10-46) kept in HBM between passes
int b2pt_film_create(uint32_t width, uint32_t height, b2pt_film_t* out) {
  using namespace b2pt;
  B2_CHECK(out && width > 0 && height > 0, B2PT_ERR_INVALID, "b2pt_film_create: empty film");
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); B2_CHECK(false, B2PT_ERR_NO_DEVICE, "b2pt_film_create: no CUDA device (there is no CPU fallback)"); }
  b2pt_film_s* f = new b2pt_film_s();
  f->width = width; f->height = height;
  cudaError_t e = cudaGetDevice(&f->device);
  if (e == cudaSuccess) e = cudaMalloc(&f->data, size_t(width) * height * 4 * sizeof(double));
  if (e == cudaSuccess) e = cudaMemset(f->data, 0, size_t(width) * height * 4 * sizeof(double));
  if (e != cudaSuccess) { if (f->data) cudaFree(f->data); delete f; return cuda_fail(e, "b2pt_film_create", __FILE__, __LINE__); }
  *out = f;
  return B2PT_OK;
}

void b2pt_film_destroy(b2pt_film_t film) {
  if (!film) return;
  cudaFree(film->data);
  delete film;
}

double* b2pt_film_data(b2pt_film_t film) { return film ? film->data : nullptr; }

int b2pt_film_clear(b2pt_film_t film, void* stream) {
  using namespace b2pt;
  B2_CHECK(film, B2PT_ERR_INVALID, "b2pt_film_clear: null film");
  B2_CUDA(cudaMemsetAsync(film->data, 0, size_t(film->width) * film->height * 4 * sizeof(double), (cudaStream_t)stream));
  return B2PT_OK;
}

int b2pt_film_read(b2pt_film_t film, double* view_host, int add, void* stream) {
  using namespace b2pt;
  B2_CHECK(film && view_host, B2PT_ERR_INVALID, "b2pt_film_read: null argument");
  const size_t n = size_t(film->width) * film->height * 4;
  if (!add) {
    B2_CUDA(cudaMemcpyAsync(view_host, film->data, n * sizeof(double), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    B2_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return B2PT_OK;
  }
  double* tmp = nullptr;
  B2_CUDA(cudaMallocHost((void**)&tmp, n * sizeof(double)));
  cudaError_t e = cudaMemcpyAsync(tmp, film->data, n * sizeof(double), cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);
  if (e == cudaSuccess) for (size_t i = 0; i < n; ++i) view_host[i] += tmp[i];
  cudaFreeHost(tmp);
  B2_CUDA(e);
  return B2PT_OK;
}

}  // extern "C"
