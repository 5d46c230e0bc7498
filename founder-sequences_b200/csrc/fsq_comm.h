// Communicators of the column-sharded long path (internal).  The one collective the path needs is an all-gather of
// equally sized device buffers, enqueued on the caller's stream:
//   * nccl_comm   one rank per GPU, ncclAllGather over NVLink / NVSwitch (libnccl is opened at run time: the library
//                 itself has no link-time dependency on it)
//   * local_comm  ranks as threads of one process, possibly on one device (the 1-GPU test boxes): device copies ordered
//                 by events, the hosts meet at a barrier
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
#include <string>

struct fsq_comm
{
	int rank{0}, size{1};
	std::string error;
	virtual ~fsq_comm() {}
	// recv holds size blocks of `bytes`; block r comes from rank r's send.  In place when send == recv + rank * bytes.
	virtual cudaError_t all_gather(void const *send, void *recv, size_t bytes, cudaStream_t st) = 0;
	// Several all-gathers launched as one (ncclGroupStart / ncclGroupEnd); a no-op for the local communicator.
	virtual void group_begin() {}
	virtual cudaError_t group_end() { return cudaSuccess; }
};

struct fsq_local_group;

namespace fsq {

fsq_local_group *local_group_create(int size);
void local_group_free(fsq_local_group *g);
fsq_comm *local_comm_create(fsq_local_group *g, int rank);

// id: 128 bytes (ncclUniqueId).  Both return a message in *err on failure.
bool nccl_unique_id(uint8_t *id, std::string *err);
fsq_comm *nccl_comm_create(uint8_t const *id, int rank, int size, std::string *err);
fsq_comm *nccl_comm_adopt(void *nccl_comm, int rank, int size, std::string *err);

} // namespace fsq
