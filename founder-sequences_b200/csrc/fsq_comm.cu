// Communicators of the column-sharded long path: NCCL (one rank per GPU) and ranks-as-threads (tests on one GPU).
#include "fsq_comm.h"
#include <nccl.h>
#include <dlfcn.h>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <vector>

namespace {

// ---- NCCL, resolved at run time ------------------------------------------------------------------------------------
struct nccl_api
{
	void *handle{nullptr};
	ncclResult_t (*GetUniqueId)(ncclUniqueId *){};
	ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int){};
	ncclResult_t (*CommDestroy)(ncclComm_t){};
	ncclResult_t (*AllGather)(void const *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t){};
	ncclResult_t (*GroupStart)(){};
	ncclResult_t (*GroupEnd)(){};
	char const *(*GetErrorString)(ncclResult_t){};
	std::string error;
};

nccl_api *load_nccl()
{
	static nccl_api api;
	static std::once_flag once;
	std::call_once(once, [] {
		// a libnccl the process has loaded already (torch brings its own) is found first by its soname
		char const *names[] = {"libnccl.so.2", "libnccl.so", nullptr};
		for (char const **n = names; *n && !api.handle; ++n) api.handle = dlopen(*n, RTLD_NOW | RTLD_GLOBAL);
		if (!api.handle) { api.error = std::string("libnccl.so.2 cannot be loaded: ") + dlerror(); return; }
		auto sym = [&](char const *name) -> void * {
			void *p = dlsym(api.handle, name);
			if (!p && api.error.empty()) api.error = std::string("libnccl lacks ") + name;
			return p;
		};
		api.GetUniqueId = (decltype(api.GetUniqueId)) sym("ncclGetUniqueId");
		api.CommInitRank = (decltype(api.CommInitRank)) sym("ncclCommInitRank");
		api.CommDestroy = (decltype(api.CommDestroy)) sym("ncclCommDestroy");
		api.AllGather = (decltype(api.AllGather)) sym("ncclAllGather");
		api.GroupStart = (decltype(api.GroupStart)) sym("ncclGroupStart");
		api.GroupEnd = (decltype(api.GroupEnd)) sym("ncclGroupEnd");
		api.GetErrorString = (decltype(api.GetErrorString)) sym("ncclGetErrorString");
	});
	return &api;
}

struct nccl_comm final : fsq_comm
{
	nccl_api *api{};
	ncclComm_t comm{};
	bool owned{true};
	ncclResult_t first_error{ncclSuccess};

	~nccl_comm() override { if (owned && comm) api->CommDestroy(comm); }

	cudaError_t note(ncclResult_t r)
	{
		if (ncclSuccess == r) return cudaSuccess;
		if (ncclSuccess == first_error) { first_error = r; error = std::string("NCCL: ") + api->GetErrorString(r); }
		return cudaErrorUnknown;
	}

	cudaError_t all_gather(void const *send, void *recv, size_t bytes, cudaStream_t st) override
	{
		return note(api->AllGather(send, recv, bytes, ncclUint8, comm, st));
	}
	void group_begin() override { note(api->GroupStart()); }
	cudaError_t group_end() override { return note(api->GroupEnd()); }
};

} // namespace

// ---- ranks as threads ------------------------------------------------------------------------------------------------
// The hosts meet at a barrier; on the device side rank r's stream waits for every other rank's "send is ready" event,
// copies, and the senders' streams wait for every reader's "copied" event before they go on (a sender may overwrite
// its buffer right after the collective).
struct fsq_local_group
{
	int size{};
	std::mutex mu;
	std::condition_variable cv;
	int waiting{0};
	uint64_t generation{0};
	std::vector <void const *> send;
	std::vector <cudaEvent_t> ready, copied;
	std::vector <int> device;

	bool broken{false};

	// false when a rank failed to arrive within two minutes (it returned early with an error): the group is then broken
	// for good and every later collective fails at once instead of waiting
	bool barrier()
	{
		std::unique_lock <std::mutex> lk(mu);
		if (broken) return false;
		uint64_t const gen = generation;
		if (++waiting == size) { waiting = 0; ++generation; cv.notify_all(); return true; }
		if (!cv.wait_for(lk, std::chrono::seconds(120), [&] { return gen != generation || broken; }) || broken)
		{
			broken = true;
			cv.notify_all();
			return false;
		}
		return true;
	}
};

namespace {

struct local_comm final : fsq_comm
{
	fsq_local_group *g{};

	cudaError_t all_gather(void const *send, void *recv, size_t bytes, cudaStream_t st) override
	{
		cudaError_t e = cudaSuccess;
		int dev = 0;
		cudaGetDevice(&dev);
		if (!g->ready[rank])
		{
			if (cudaSuccess != (e = cudaEventCreateWithFlags(&g->ready[rank], cudaEventDisableTiming))) return e;
			if (cudaSuccess != (e = cudaEventCreateWithFlags(&g->copied[rank], cudaEventDisableTiming))) return e;
		}
		g->send[rank] = send;
		g->device[rank] = dev;
		e = cudaEventRecord(g->ready[rank], st);
		auto broken = [&] { error = "ranks-as-threads communicator: a rank did not arrive at the collective"; return cudaErrorUnknown; };
		if (!g->barrier()) return broken();             // every rank's event is recorded and its pointer published
		for (int r = 0; r < size && cudaSuccess == e; ++r)
		{
			uint8_t *dst = static_cast <uint8_t *>(recv) + (size_t) r * bytes;
			if (r == rank)
			{
				if (dst != send) e = cudaMemcpyAsync(dst, send, bytes, cudaMemcpyDeviceToDevice, st);
				continue;
			}
			e = cudaStreamWaitEvent(st, g->ready[r], 0);
			if (cudaSuccess == e && bytes) e = cudaMemcpyPeerAsync(dst, dev, g->send[r], g->device[r], bytes, st);
		}
		if (cudaSuccess == e) e = cudaEventRecord(g->copied[rank], st);
		if (!g->barrier()) return broken();             // every rank's copies are enqueued
		for (int r = 0; r < size && cudaSuccess == e; ++r)
			if (r != rank) e = cudaStreamWaitEvent(st, g->copied[r], 0);
		if (!g->barrier()) return broken();             // nobody re-records an event another rank has yet to wait on
		return e;
	}
};

} // namespace

namespace fsq {

fsq_local_group *local_group_create(int size)
{
	if (size < 1) return nullptr;
	fsq_local_group *g = new fsq_local_group;
	g->size = size;
	g->send.assign(size, nullptr);
	g->ready.assign(size, nullptr);
	g->copied.assign(size, nullptr);
	g->device.assign(size, 0);
	return g;
}

void local_group_free(fsq_local_group *g)
{
	if (!g) return;
	for (auto ev : g->ready) if (ev) cudaEventDestroy(ev);
	for (auto ev : g->copied) if (ev) cudaEventDestroy(ev);
	delete g;
}

fsq_comm *local_comm_create(fsq_local_group *g, int rank)
{
	if (!g || rank < 0 || rank >= g->size) return nullptr;
	local_comm *c = new local_comm;
	c->g = g; c->rank = rank; c->size = g->size;
	return c;
}

bool nccl_unique_id(uint8_t *id, std::string *err)
{
	nccl_api *api = load_nccl();
	if (!api->error.empty()) { *err = api->error; return false; }
	static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
	ncclUniqueId u;
	ncclResult_t const r = api->GetUniqueId(&u);
	if (ncclSuccess != r) { *err = std::string("ncclGetUniqueId: ") + api->GetErrorString(r); return false; }
	memcpy(id, &u, sizeof u);
	return true;
}

fsq_comm *nccl_comm_create(uint8_t const *id, int rank, int size, std::string *err)
{
	nccl_api *api = load_nccl();
	if (!api->error.empty()) { *err = api->error; return nullptr; }
	ncclUniqueId u;
	memcpy(&u, id, sizeof u);
	nccl_comm *c = new nccl_comm;
	c->api = api; c->rank = rank; c->size = size;
	ncclResult_t const r = api->CommInitRank(&c->comm, size, u, rank);
	if (ncclSuccess != r)
	{
		*err = std::string("ncclCommInitRank: ") + api->GetErrorString(r);
		c->comm = nullptr;
		delete c;
		return nullptr;
	}
	// NCCL connects its peers at the first collective (0.5 - 3 s at 2 - 8 GPUs, measured as the first call's time): one small
	// all-gather here, where every rank is anyway, so that the first fsq_segment_sharded finds the connections made
	{
		void *buf = nullptr;
		cudaError_t e = cudaMalloc(&buf, (size_t) size * 64);
		if (cudaSuccess == e) e = cudaMemset(buf, 0, (size_t) size * 64);
		if (cudaSuccess == e) e = c->all_gather(static_cast <uint8_t *>(buf) + (size_t) rank * 64, buf, 64, nullptr);
		if (cudaSuccess == e) e = cudaStreamSynchronize(nullptr);
		if (buf) cudaFree(buf);
		if (cudaSuccess != e)
		{
			*err = c->error.empty() ? std::string("first all-gather of the new communicator: ") + cudaGetErrorString(e) : c->error;
			delete c;
			return nullptr;
		}
	}
	return c;
}

fsq_comm *nccl_comm_adopt(void *comm, int rank, int size, std::string *err)
{
	nccl_api *api = load_nccl();
	if (!api->error.empty()) { *err = api->error; return nullptr; }
	nccl_comm *c = new nccl_comm;
	c->api = api; c->rank = rank; c->size = size; c->comm = static_cast <ncclComm_t>(comm); c->owned = false;
	return c;
}

} // namespace fsq
