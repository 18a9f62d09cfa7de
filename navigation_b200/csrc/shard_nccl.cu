// shard_nccl.cu -- collectives of the sharded filter on an NCCL communicator the
// library owns (NCCL call sites next to the kernels).
//
// The reference has no communication layer; a sharded filter needs a handful of
// exchanges per update (weight totals, the cumulative table, histogram merge,
// per-cluster sums).  amclb200_pf_set_shard() takes them as callbacks from the
// host program, which costs a trip through the caller (Python: ~50 us each) for
// every one of them.  With amclb200_pf_set_shard_nccl() the library creates its
// own communicator from a 128-byte ncclUniqueId the host program distributes
// once, and enqueues ncclAllReduce / ncclAllGather on its stream itself; a C++
// amcl_node needs nothing but MPI-style rank / world / id plumbing.
//
// libnccl is opened with dlopen on first use, so a single-GPU user of
// libamcl_b200.so does not need NCCL installed; inside a PyTorch process the
// libnccl.so.2 torch has already loaded is the one that answers.
#include <dlfcn.h>
#include <string.h>

#include "context.hpp"

namespace amclb200
{

namespace
{

// The few declarations of nccl.h this file needs (stable since NCCL 2.0).
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclInt8 = 0, ncclInt32 = 2, ncclInt64 = 4, ncclFloat64 = 8 };
enum { ncclSum = 0, ncclProd = 1, ncclMax = 2, ncclMin = 3 };

struct Api
{
  int (*GetUniqueId)(ncclUniqueId *) = nullptr;
  int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
  bool ready = false;
};

Api &api()
{
  static Api a;
  if (a.ready) return a;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) AMCL_FATAL("multi-GPU filter: cannot open libnccl.so.2 (%s)", dlerror());
#define AMCL_NCCL_SYM(field, name)                                                   \
  do                                                                                 \
  {                                                                                  \
    *(void **)(&a.field) = dlsym(h, name);                                           \
    if (!a.field) AMCL_FATAL("multi-GPU filter: %s is missing from libnccl", name);  \
  } while (0)
  AMCL_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
  AMCL_NCCL_SYM(CommInitRank, "ncclCommInitRank");
  AMCL_NCCL_SYM(CommDestroy, "ncclCommDestroy");
  AMCL_NCCL_SYM(AllReduce, "ncclAllReduce");
  AMCL_NCCL_SYM(AllGather, "ncclAllGather");
  AMCL_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef AMCL_NCCL_SYM
  a.ready = true;
  return a;
}

#define AMCL_NCCL(call)                                                                  \
  do                                                                                     \
  {                                                                                      \
    int rc__ = (call);                                                                   \
    if (rc__ != ncclSuccess)                                                             \
      AMCL_FATAL("NCCL failure %s in `%s`", api().GetErrorString(rc__), #call);          \
  } while (0)

}  // namespace

void nccl_allreduce(void *comm, void *d_buf, int count, int dtype, int op)
{
  if (count <= 0) return;
  const int dt = dtype == SHARD_F64 ? ncclFloat64 : (dtype == SHARD_I64 ? ncclInt64 : ncclInt32);
  const int o = op == SHARD_SUM ? ncclSum : (op == SHARD_MIN ? ncclMin : ncclMax);
  AMCL_NCCL(api().AllReduce(d_buf, d_buf, (size_t)count, dt, o, (ncclComm_t)comm, globals().stream));
}

// In place: rank r's block already sits at r * bytes_per_rank.
void nccl_allgather(void *comm, int rank, void *d_buf, unsigned long long bytes_per_rank)
{
  if (bytes_per_rank == 0) return;
  const char *mine = (const char *)d_buf + (size_t)rank * bytes_per_rank;
  AMCL_NCCL(api().AllGather(mine, d_buf, (size_t)bytes_per_rank, ncclInt8, (ncclComm_t)comm,
                            globals().stream));
}

void nccl_release(void *comm)
{
  if (comm) api().CommDestroy((ncclComm_t)comm);
}

void Shard::allreduce(void *d_buf, int count, int dtype, int op, void *) const
{
  if (nccl_comm) nccl_allreduce(nccl_comm, d_buf, count, dtype, op);
  else cb_allreduce(d_buf, count, dtype, op, user);
}

void Shard::allgather(void *d_buf, unsigned long long bytes_per_rank, void *) const
{
  if (nccl_comm) nccl_allgather(nccl_comm, rank, d_buf, bytes_per_rank);
  else cb_allgather(d_buf, bytes_per_rank, user);
}

}  // namespace amclb200

using namespace amclb200;

extern "C" {

// Rank 0 calls this and hands the 128 bytes to every other rank (any host-side
// channel: MPI_Bcast, torch.distributed.broadcast, a file).
int amclb200_nccl_unique_id(unsigned char out[128])
{
  ncclUniqueId id;
  const int rc = api().GetUniqueId(&id);
  if (rc != ncclSuccess) return -1;
  memcpy(out, id.internal, 128);
  return 0;
}

// Collective: every rank calls it with the same id.  From here on the filter is
// rank `rank` of `world` exactly as after amclb200_pf_set_shard(), with the
// exchanges running on a communicator owned by the library.
void amclb200_pf_set_shard_nccl(pf_t *pf, int rank, int world, const unsigned char id_bytes[128])
{
  PfCtx *c = pf_ctx(pf);
  if (world < 1 || rank < 0 || rank >= world || world > 1024)
    AMCL_FATAL("amclb200_pf_set_shard_nccl: rank %d of %d", rank, world);
  ensure_device();
  c->shard.rank = rank;
  c->shard.world = world;
  c->shard.align = world > 1 && world <= kShardMaxAligned ? kShardTile : 1;
  c->shard.cb_allreduce = nullptr;
  c->shard.cb_allgather = nullptr;
  c->shard.user = nullptr;
  if (c->shard.nccl_comm) nccl_release(c->shard.nccl_comm);
  c->shard.nccl_comm = nullptr;
  if (world == 1) return;
  ncclUniqueId id;
  memcpy(id.internal, id_bytes, 128);
  ncclComm_t comm = nullptr;
  AMCL_NCCL(api().CommInitRank(&comm, world, id, rank));
  c->shard.nccl_comm = comm;
}

}  // extern "C"
