// NCCL over NVLink/NVSwitch for the subtree-partitioned solver (kernel row K9).
//
// One process per GPU.  The library does not link NCCL: libnccl.so.2 is opened at run time the first
// time a communicator is asked for, so a single-GPU user never needs it and, inside a process that
// already loaded torch's NCCL, the very same copy is used.  Only point-to-point transfers of
// contribution blocks / boundary vectors and two small all-reduces (inertia counters, solution
// vector) go through it.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>

#include <cuda_runtime.h>

namespace b200kkt {

constexpr int COMM_ID_BYTES = 128;   // NCCL_UNIQUE_ID_BYTES

struct Comm {
   void* nccl = nullptr;   // ncclComm_t
   int rank = 0, world = 1;

   bool active() const { return nccl != nullptr; }
};

// all return 0 on success; on failure `err` holds the NCCL / loader message
int comm_unique_id(char id[COMM_ID_BYTES], std::string& err);
int comm_init(Comm& c, int rank, int world, const char id[COMM_ID_BYTES], std::string& err);
void comm_destroy(Comm& c);

int comm_group_start(std::string& err);
int comm_group_end(std::string& err);
int comm_send_f64(const Comm& c, const double* p, size_t count, int peer, cudaStream_t st, std::string& err);
int comm_recv_f64(const Comm& c, double* p, size_t count, int peer, cudaStream_t st, std::string& err);
int comm_bcast_bytes(const Comm& c, void* p, size_t bytes, int root, cudaStream_t st, std::string& err);   // in place
int comm_allreduce_sum_f64(const Comm& c, double* p, size_t count, cudaStream_t st, std::string& err);
int comm_allreduce_sum_i32(const Comm& c, int* p, size_t count, cudaStream_t st, std::string& err);

}  // namespace b200kkt
