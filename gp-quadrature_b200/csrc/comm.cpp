// The library's own NCCL communicator: the two exchanges of the point-sharded fit (MAX of the bounding box, SUM of
// the two cropped mode grids) are issued from C on the stream of the kernels that produce
// their inputs, with no Python between the spread epilogue and the reduction.
//
// NCCL is resolved with dlopen at communicator creation ("libnccl.so.2": the copy PyTorch already loaded when the host
// program uses torch.distributed, else the system library), so the library has no link-time dependency and single-GPU
// users never touch it.  Only the stable core API is used; the declarations below restate nccl.h for those calls.
#include "common.cuh"
#include <dlfcn.h>
#include <mutex>

namespace efgp {

namespace {
struct NcclId { char internal[128]; };             // ncclUniqueId (NCCL_UNIQUE_ID_BYTES = 128)
typedef void* NcclComm;                            // ncclComm_t
typedef int (*fn_GetUniqueId)(NcclId*);
typedef int (*fn_CommInitRank)(NcclComm*, int, NcclId, int);
typedef int (*fn_CommDestroy)(NcclComm);
typedef int (*fn_AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef const char* (*fn_GetErrorString)(int);
constexpr int kNcclSum = 0, kNcclMax = 2, kNcclFloat32 = 7, kNcclFloat64 = 8;

struct Api {
  void* lib = nullptr;
  fn_GetUniqueId GetUniqueId = nullptr;
  fn_CommInitRank CommInitRank = nullptr;
  fn_CommDestroy CommDestroy = nullptr;
  fn_AllReduce AllReduce = nullptr;
  fn_GetErrorString GetErrorString = nullptr;
};
Api g_api;
NcclComm g_comm = nullptr;
int g_rank = 0, g_world = 1;
std::mutex g_cmu;

void load_api() {
  if (g_api.lib) return;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    g_api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_api.lib) break;
  }
  if (!g_api.lib) throw Error(5, std::string("NCCL not found (dlopen libnccl.so.2): ") + dlerror());
  g_api.GetUniqueId = (fn_GetUniqueId)dlsym(g_api.lib, "ncclGetUniqueId");
  g_api.CommInitRank = (fn_CommInitRank)dlsym(g_api.lib, "ncclCommInitRank");
  g_api.CommDestroy = (fn_CommDestroy)dlsym(g_api.lib, "ncclCommDestroy");
  g_api.AllReduce = (fn_AllReduce)dlsym(g_api.lib, "ncclAllReduce");
  g_api.GetErrorString = (fn_GetErrorString)dlsym(g_api.lib, "ncclGetErrorString");
  if (!g_api.GetUniqueId || !g_api.CommInitRank || !g_api.CommDestroy || !g_api.AllReduce)
    throw Error(5, "NCCL library lacks the core API symbols");
}

void nccl_check(int r, const char* what) {
  if (r != 0)
    throw Error(5, std::string(what) + ": NCCL error " + std::to_string(r) + " (" +
                       (g_api.GetErrorString ? g_api.GetErrorString(r) : "?") + ")");
}
}  // namespace

void comm_unique_id(void* id128) {
  std::lock_guard<std::mutex> lk(g_cmu);
  load_api();
  NcclId id;
  nccl_check(g_api.GetUniqueId(&id), "ncclGetUniqueId");
  memcpy(id128, id.internal, 128);
}

void comm_init(int rank, int world, const void* id128) {
  std::lock_guard<std::mutex> lk(g_cmu);
  EFGP_CHECK(world >= 1 && rank >= 0 && rank < world, "bad rank / world size");
  load_api();
  if (g_comm) {
    g_api.CommDestroy(g_comm);
    g_comm = nullptr;
  }
  NcclId id;
  memcpy(id.internal, id128, 128);
  nccl_check(g_api.CommInitRank(&g_comm, world, id, rank), "ncclCommInitRank");
  g_rank = rank;
  g_world = world;
}

void comm_destroy() {
  std::lock_guard<std::mutex> lk(g_cmu);
  if (g_comm && g_api.CommDestroy) g_api.CommDestroy(g_comm);
  g_comm = nullptr;
  g_world = 1;
  g_rank = 0;
}

int comm_world() { return g_comm ? g_world : 1; }
int comm_rank() { return g_comm ? g_rank : 0; }

// in place; dtype 0 = float32, 1 = float64; op 0 = sum, 1 = max.  Enqueues on st.
void comm_allreduce(void* buf, int64_t count, int dtype, int op, cudaStream_t st) {
  if (!g_comm || g_world == 1) return;
  EFGP_CHECK(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
  EFGP_CHECK(op == 0 || op == 1, "op must be 0 (sum) or 1 (max)");
  nccl_check(g_api.AllReduce(buf, buf, (size_t)count, dtype == 0 ? kNcclFloat32 : kNcclFloat64,
                             op == 0 ? kNcclSum : kNcclMax, g_comm, st),
             "ncclAllReduce");
}

}  // namespace efgp
