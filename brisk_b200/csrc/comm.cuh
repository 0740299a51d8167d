// comm.cuh — the exchange step of the sharded index: one process per GPU, the 4^b bucket ids
// split over the ranks, superkmer records exchanged all-to-all with NCCL over NVLink / NVSwitch so that every
// rank inserts (or looks up) only the buckets it owns.  Replaces the lock-stripe routing of
// DenseMenuYo::insert_kmer_vector / get_kmer_vector (brisk/DenseMenuYo.hpp:248-308, :352-396) across GPUs.
//
// NCCL is bound at run time (dlopen): inside a torch process the already-loaded libnccl.so.2 is reused, so the
// library and torch.distributed share one NCCL; a plain C++ host gets the system library.  A single-GPU index
// never touches NCCL.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <string>
#include <vector>

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;

    bool load() {
        if (lib) return true;
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {  // an NCCL that is already mapped (torch's) wins
            lib = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
            if (lib) break;
        }
        for (int i = 0; i < 2 && !lib; i++) lib = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
        if (!lib) {
            error = std::string("cannot load libnccl: ") + dlerror();
            return false;
        }
#define NCCL_SYM(field, name)                                        \
    *(void **)(&field) = dlsym(lib, name);                           \
    if (!field) { error = std::string("libnccl lacks ") + name; lib = nullptr; return false; }
        NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
        NCCL_SYM(CommInitRank, "ncclCommInitRank")
        NCCL_SYM(CommDestroy, "ncclCommDestroy")
        NCCL_SYM(Send, "ncclSend")
        NCCL_SYM(Recv, "ncclRecv")
        NCCL_SYM(AllGather, "ncclAllGather")
        NCCL_SYM(AllReduce, "ncclAllReduce")
        NCCL_SYM(GroupStart, "ncclGroupStart")
        NCCL_SYM(GroupEnd, "ncclGroupEnd")
        NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef NCCL_SYM
        return true;
    }
};

static NcclApi g_nccl;

// the send / receive layout of one exchange, from the n x n matrix of record counts (row = sender):
// rank r sends counts[r][d] records to d from offset send_start[d] of its partitioned buffer and receives
// counts[s][r] records from s at offset recv_start[s] of its receive buffer.  Pure host arithmetic, exported
// through the C ABI (brisk_exchange_plan) so that the CPU tests can check it.
static inline void exchange_plan(const uint64_t *counts, int n, int rank, uint64_t *send_start, uint64_t *recv_start) {
    uint64_t acc = 0;
    for (int d = 0; d < n; d++) {
        send_start[d] = acc;
        acc += counts[(size_t)rank * n + d];
    }
    send_start[n] = acc;
    acc = 0;
    for (int s = 0; s < n; s++) {
        recv_start[s] = acc;
        acc += counts[(size_t)s * n + rank];
    }
    recv_start[n] = acc;
}

struct Comm {
    ncclComm_t comm = nullptr;
    int rank = 0, n = 1;
    // The exchange runs on its own stream: while the records of chunk c travel, the handle's stream enumerates chunk
    // c + 1 (brisk_capi.cu: exchange_pipelined).  The host only ever waits on this stream's events.
    cudaStream_t stream = nullptr;
    cudaStream_t aux[2] = {nullptr, nullptr};  // the enumerate chunks before the last one (decreasing priority, overlapping launch tails)
    bool aux_ordered = false;                  // the device has enough priority levels for that
    cudaEvent_t ev_enum[16] = {}, ev_counts = nullptr, ev_done = nullptr;
    uint64_t *h_pinned = nullptr;  // pinned: [0, 16) record counters after every chunk, [16, 16 + 65 * 64) gathered totals + status
    u64 *d_gather = nullptr;       // device: partition scratch whose first n + 1 words are this rank's totals + status, then [64][65] everybody's
    std::vector<uint64_t> h_matrix, send_start, recv_start;
};
#define COMM_MAX_CHUNKS 16
