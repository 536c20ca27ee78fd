// K14: gradient all-reduce over NCCL / NVLink (nothing in the reference; the
// data-parallel replicas of north_star item 4).  NCCL is resolved at run time
// with dlopen("libnccl.so.2") - the copy torch already loaded in this process -
// so liboflow_b200.so has no link-time NCCL dependency and single-GPU use never
// touches it.
#include <dlfcn.h>
#include "common.cuh"

typedef struct { char internal[128]; } of_nccl_uid;
typedef void* of_nccl_comm;
typedef int (*fn_get_uid)(of_nccl_uid*);
typedef int (*fn_init_rank)(of_nccl_comm*, int, of_nccl_uid, int);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, of_nccl_comm, cudaStream_t);
typedef int (*fn_destroy)(of_nccl_comm);
typedef const char* (*fn_errstr)(int);

static struct {
  void* h;
  fn_get_uid get_uid;
  fn_init_rank init_rank;
  fn_allreduce allreduce;
  fn_destroy destroy;
  fn_errstr errstr;
} g_nccl;

static int load_nccl() {
  if (g_nccl.h) return 0;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  OF_REQUIRE(h != nullptr, OF_ERR_STATE, "dlopen(libnccl.so.2) failed: %s", dlerror());
  g_nccl.get_uid = (fn_get_uid)dlsym(h, "ncclGetUniqueId");
  g_nccl.init_rank = (fn_init_rank)dlsym(h, "ncclCommInitRank");
  g_nccl.allreduce = (fn_allreduce)dlsym(h, "ncclAllReduce");
  g_nccl.destroy = (fn_destroy)dlsym(h, "ncclCommDestroy");
  g_nccl.errstr = (fn_errstr)dlsym(h, "ncclGetErrorString");
  OF_REQUIRE(g_nccl.get_uid && g_nccl.init_rank && g_nccl.allreduce && g_nccl.destroy && g_nccl.errstr,
             OF_ERR_STATE, "libnccl.so.2 lacks a required symbol");
  g_nccl.h = h;
  return 0;
}

#define OF_NCCL(expr)                                                              \
  do {                                                                             \
    int _r = (expr);                                                               \
    if (_r != 0) {                                                                 \
      of_set_error("%s failed: %s", #expr, g_nccl.errstr ? g_nccl.errstr(_r) : "?"); \
      return OF_ERR_STATE;                                                         \
    }                                                                              \
  } while (0)

extern "C" {

int of_comm_unique_id(void* host_id128) {
  OF_REQUIRE(host_id128, OF_ERR_ARG, "null pointer argument");
  if (int r = load_nccl()) return r;
  of_nccl_uid id;
  OF_NCCL(g_nccl.get_uid(&id));
  memcpy(host_id128, &id, sizeof(id));
  return 0;
}

int of_comm_init(const void* host_id128, int world_size, int rank, void** comm_out) {
  OF_REQUIRE(host_id128 && comm_out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(world_size >= 1 && rank >= 0 && rank < world_size, OF_ERR_ARG, "bad rank %d / world %d", rank,
             world_size);
  if (int r = load_nccl()) return r;
  of_nccl_uid id;
  memcpy(&id, host_id128, sizeof(id));
  of_nccl_comm c = nullptr;
  OF_NCCL(g_nccl.init_rank(&c, world_size, id, rank));
  *comm_out = c;
  return 0;
}

int of_comm_allreduce_sum_f32(void* comm, float* buf, size_t n, void* stream) {
  OF_REQUIRE(comm && buf, OF_ERR_ARG, "null pointer argument");
  // ncclFloat32 = 7, ncclSum = 0
  OF_NCCL(g_nccl.allreduce(buf, buf, n, 7, 0, (of_nccl_comm)comm, (cudaStream_t)stream));
  return 0;
}

int of_comm_destroy(void* comm) {
  if (!comm) return 0;
  if (int r = load_nccl()) return r;
  OF_NCCL(g_nccl.destroy((of_nccl_comm)comm));
  return 0;
}

}  // extern "C"
