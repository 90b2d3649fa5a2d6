// Run-time binding to NCCL (see comm.hpp).  Only the handful of entry points the partitioned solver
// needs; types are the public ones of nccl.h (opaque communicator, 128-byte unique id, enum values
// ncclInt32 = 2, ncclFloat64 = 8, ncclSum = 0, stable since NCCL 2.0).
#include "comm.hpp"

#include <dlfcn.h>

#include <cstring>
#include <mutex>

namespace b200kkt {
namespace {

struct UniqueId { char internal[COMM_ID_BYTES]; };
using comm_t = void*;
constexpr int kInt8 = 0, kInt32 = 2, kFloat64 = 8, kSum = 0;

struct Api {
   void* lib = nullptr;
   int (*GetUniqueId)(UniqueId*) = nullptr;
   int (*CommInitRank)(comm_t*, int, UniqueId, int) = nullptr;
   int (*CommDestroy)(comm_t) = nullptr;
   const char* (*GetErrorString)(int) = nullptr;
   int (*GroupStart)() = nullptr;
   int (*GroupEnd)() = nullptr;
   int (*Send)(const void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
   int (*Recv)(void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
   int (*AllReduce)(const void*, void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
   int (*Broadcast)(const void*, void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
   std::string load_error;
};

Api& api() {
   static Api a;
   static std::once_flag once;
   std::call_once(once, [] {
      for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
         a.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
         if (a.lib != nullptr) break;
      }
      if (a.lib == nullptr) {
         const char* e = dlerror();
         a.load_error = std::string("cannot open libnccl.so.2: ") + (e != nullptr ? e : "unknown loader error");
         return;
      }
      bool ok = true;
      auto sym = [&](const char* n) {
         void* p = dlsym(a.lib, n);
         if (p == nullptr) { ok = false; a.load_error = std::string("libnccl lacks ") + n; }
         return p;
      };
      a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
      a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
      a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
      a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
      a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
      a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
      a.Send = reinterpret_cast<decltype(a.Send)>(sym("ncclSend"));
      a.Recv = reinterpret_cast<decltype(a.Recv)>(sym("ncclRecv"));
      a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(sym("ncclAllReduce"));
      a.Broadcast = reinterpret_cast<decltype(a.Broadcast)>(sym("ncclBroadcast"));
      if (!ok) { dlclose(a.lib); a.lib = nullptr; }
   });
   return a;
}

int check(int rc, const char* what, std::string& err) {
   if (rc == 0) return 0;
   Api& a = api();
   err = std::string("NCCL error in ") + what + ": " + (a.GetErrorString != nullptr ? a.GetErrorString(rc) : "?");
   return rc;
}

int ready(std::string& err) {
   Api& a = api();
   if (a.lib == nullptr) { err = a.load_error; return -1; }
   return 0;
}

}  // namespace

int comm_unique_id(char id[COMM_ID_BYTES], std::string& err) {
   if (ready(err) != 0) return -1;
   UniqueId u;
   if (int rc = check(api().GetUniqueId(&u), "ncclGetUniqueId", err)) return rc;
   std::memcpy(id, u.internal, COMM_ID_BYTES);
   return 0;
}

int comm_init(Comm& c, int rank, int world, const char id[COMM_ID_BYTES], std::string& err) {
   if (ready(err) != 0) return -1;
   comm_destroy(c);
   UniqueId u;
   std::memcpy(u.internal, id, COMM_ID_BYTES);
   comm_t comm = nullptr;
   if (int rc = check(api().CommInitRank(&comm, world, u, rank), "ncclCommInitRank", err)) return rc;
   c.nccl = comm;
   c.rank = rank;
   c.world = world;
   return 0;
}

void comm_destroy(Comm& c) {
   if (c.nccl != nullptr && api().lib != nullptr) api().CommDestroy(c.nccl);
   c.nccl = nullptr;
}

int comm_group_start(std::string& err) { return check(api().GroupStart(), "ncclGroupStart", err); }
int comm_group_end(std::string& err) { return check(api().GroupEnd(), "ncclGroupEnd", err); }

int comm_send_f64(const Comm& c, const double* p, size_t count, int peer, cudaStream_t st, std::string& err) {
   return check(api().Send(p, count, kFloat64, peer, c.nccl, st), "ncclSend", err);
}

int comm_recv_f64(const Comm& c, double* p, size_t count, int peer, cudaStream_t st, std::string& err) {
   return check(api().Recv(p, count, kFloat64, peer, c.nccl, st), "ncclRecv", err);
}

int comm_bcast_bytes(const Comm& c, void* p, size_t bytes, int root, cudaStream_t st, std::string& err) {
   // whole doubles where possible: NCCL moves the same bytes either way, the count stays small
   if (bytes % 8 == 0) return check(api().Broadcast(p, p, bytes / 8, kFloat64, root, c.nccl, st), "ncclBroadcast", err);
   return check(api().Broadcast(p, p, bytes, kInt8, root, c.nccl, st), "ncclBroadcast", err);
}

int comm_allreduce_sum_f64(const Comm& c, double* p, size_t count, cudaStream_t st, std::string& err) {
   return check(api().AllReduce(p, p, count, kFloat64, kSum, c.nccl, st), "ncclAllReduce", err);
}

int comm_allreduce_sum_i32(const Comm& c, int* p, size_t count, cudaStream_t st, std::string& err) {
   return check(api().AllReduce(p, p, count, kInt32, kSum, c.nccl, st), "ncclAllReduce", err);
}

}  // namespace b200kkt
