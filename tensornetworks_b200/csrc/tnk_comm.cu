// tnk_comm: the exchange step of the sharded evaluations behind the C-ABI.
//
// One process per GPU on one NVSwitch box.  Every rank owns an exchange buffer in device memory
//     [ send : cap doubles | recv : cap doubles | flags : TNK_MAX_PEERS x uint64 ]
// exported as a CUDA IPC handle; the host all-gathers the 64-byte handles (any transport) and every rank maps
// all peers' buffers (NVLink P2P).  A sum-all-reduce of `count` doubles is then three launches of our own:
//     barrier (flag epoch)  ->  reduce-scatter + all-gather in ONE kernel  ->  barrier
// The exchange kernel on rank g owns slice g of the vector: it loads that slice from the send buffer of EVERY rank
// over NVLink, adds the G values in rank order (so all ranks end with bit-identical sums, whatever G is) and stores
// the sum into the recv buffer of EVERY rank.  The evaluation kernels write their partial results straight into
// the send buffer (tnk_comm_sendbuf), so nothing is copied before the exchange; no NCCL call is involved.
#include "tnk_common.cuh"

struct tnk_comm {
    int rank, world, device;
    int64_t cap;                       // doubles per send / recv region
    unsigned char* local;              // this rank's buffer (cudaMalloc)
    unsigned char* peer[TNK_MAX_PEERS];// all ranks' buffers as mapped here (peer[rank] == local)
    bool connected;
    uint64_t epoch;                    // barrier generation (host side; the same sequence on every rank)
};

namespace {

struct CommPtrs {
    int rank, world;
    double* send[TNK_MAX_PEERS];
    double* recv[TNK_MAX_PEERS];
    unsigned long long* flags[TNK_MAX_PEERS];
};

size_t comm_bytes(int64_t cap) { return (size_t)cap * 16 + TNK_MAX_PEERS * sizeof(uint64_t) * 2; }

CommPtrs comm_ptrs(const tnk_comm* c) {
    CommPtrs p;
    p.rank = c->rank;
    p.world = c->world;
    for (int r = 0; r < TNK_MAX_PEERS; ++r) {
        unsigned char* base = r < c->world ? c->peer[r] : nullptr;
        p.send[r] = (double*)base;
        p.recv[r] = base ? (double*)(base + (size_t)c->cap * 8) : nullptr;
        p.flags[r] = base ? (unsigned long long*)(base + (size_t)c->cap * 16) : nullptr;
    }
    return p;
}

// Cross-rank barrier: thread p publishes `epoch` in slot `rank` of peer p's flag array and waits until slot p of
// the local array has reached it.  Everything the earlier kernels of this stream wrote (also into peer memory) is
// ordered before the release store by the kernel boundary plus the system-scope fence.
__global__ void comm_barrier_kernel(const CommPtrs c, unsigned long long epoch) {
    const int p = threadIdx.x;
    if (p >= c.world) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(c.flags[p] + c.rank), "l"(epoch) : "memory");
    unsigned long long seen;
    do {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(c.flags[c.rank] + p) : "memory");
    } while (seen < epoch);
    __threadfence_system();
}

// reduce-scatter + all-gather: this rank sums slice `rank` of all send buffers (rank order) and stores it everywhere
__global__ void __launch_bounds__(256)
comm_exchange_kernel(const CommPtrs c, int64_t count) {
    const int64_t pairs = (count + 1) / 2;                       // work in double2 (the regions are 16-byte aligned)
    const int64_t lo = pairs * c.rank / c.world, hi = pairs * (c.rank + 1) / c.world;
    for (int64_t v = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < hi; v += (int64_t)gridDim.x * blockDim.x) {
        double2 acc = reinterpret_cast<const double2*>(c.send[0])[v];
        for (int p = 1; p < c.world; ++p) {
            const double2 x = reinterpret_cast<const double2*>(c.send[p])[v];
            acc.x += x.x;
            acc.y += x.y;
        }
        for (int p = 0; p < c.world; ++p) reinterpret_cast<double2*>(c.recv[p])[v] = acc;
    }
}

}  // namespace

extern "C" int tnk_comm_create(int rank, int world, int64_t max_doubles, tnk_comm** comm, unsigned char* handle_out) {
    TNK_REQUIRE(comm && handle_out, "tnk_comm_create: NULL argument");
    TNK_REQUIRE(world >= 1 && world <= TNK_MAX_PEERS && rank >= 0 && rank < world,
                "tnk_comm_create: rank %d / world %d outside [1,%d]", rank, world, TNK_MAX_PEERS);
    TNK_REQUIRE(max_doubles > 0, "tnk_comm_create: max_doubles must be positive");
    static_assert(sizeof(cudaIpcMemHandle_t) == TNK_COMM_HANDLE_BYTES, "IPC handle size");
    tnk_comm* c = new tnk_comm();
    c->rank = rank;
    c->world = world;
    c->cap = (max_doubles + 31) / 32 * 32;
    c->connected = false;
    c->epoch = 0;
    for (int r = 0; r < TNK_MAX_PEERS; ++r) c->peer[r] = nullptr;
    cudaError_t e = cudaGetDevice(&c->device);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->local, comm_bytes(c->cap));
    if (e == cudaSuccess) e = cudaMemset(c->local, 0, comm_bytes(c->cap));
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, c->local);
    if (e != cudaSuccess) {
        tnk_set_error("tnk_comm_create: %s", cudaGetErrorString(e));
        delete c;
        return 1;
    }
    memcpy(handle_out, &h, sizeof(h));
    c->peer[rank] = c->local;
    *comm = c;
    return 0;
}

extern "C" int tnk_comm_connect(tnk_comm* c, const unsigned char* all_handles) {
    TNK_REQUIRE(c && all_handles, "tnk_comm_connect: NULL argument");
    TNK_REQUIRE(!c->connected, "tnk_comm_connect: already connected");
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, all_handles + (size_t)r * TNK_COMM_HANDLE_BYTES, sizeof(h));
        void* p = nullptr;
        TNK_CUDA_OK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        c->peer[r] = (unsigned char*)p;
    }
    c->connected = true;
    return 0;
}

extern "C" int tnk_comm_destroy(tnk_comm* c) {
    if (!c) return 0;
    for (int r = 0; r < c->world; ++r)
        if (r != c->rank && c->peer[r]) cudaIpcCloseMemHandle(c->peer[r]);
    if (c->local) cudaFree(c->local);
    delete c;
    return 0;
}

extern "C" int tnk_comm_rank(const tnk_comm* c) { return c ? c->rank : -1; }
extern "C" int tnk_comm_world(const tnk_comm* c) { return c ? c->world : 0; }
extern "C" double* tnk_comm_sendbuf(tnk_comm* c) { return c ? (double*)c->local : nullptr; }
extern "C" double* tnk_comm_recvbuf(tnk_comm* c) { return c ? (double*)(c->local + (size_t)c->cap * 8) : nullptr; }

extern "C" int tnk_comm_allreduce_sum(tnk_comm* c, int64_t count, double* out, int* n_launch, void* stream) {
    TNK_REQUIRE(c, "tnk_comm_allreduce_sum: NULL communicator");
    TNK_REQUIRE(count > 0 && count <= c->cap, "tnk_comm_allreduce_sum: count %lld outside (0,%lld]", (long long)count, (long long)c->cap);
    TNK_REQUIRE(c->connected || c->world == 1, "tnk_comm_allreduce_sum: tnk_comm_connect has not been called");
    cudaStream_t st = (cudaStream_t)stream;
    int launches = 0;
    const CommPtrs p = comm_ptrs(c);
    if (c->world == 1) {
        TNK_CUDA_OK(cudaMemcpyAsync(tnk_comm_recvbuf(c), c->local, (size_t)count * 8, cudaMemcpyDeviceToDevice, st));
    } else {
        comm_barrier_kernel<<<1, 32, 0, st>>>(p, ++c->epoch);            // every rank's partial is complete
        const int64_t per_rank = ((count + 1) / 2 + c->world - 1) / c->world;
        unsigned blocks = (unsigned)((per_rank + 255) / 256);
        if (blocks > 592) blocks = 592;                                   // 4 x 148 SMs, grid-stride beyond
        if (blocks < 1) blocks = 1;
        comm_exchange_kernel<<<blocks, 256, 0, st>>>(p, count);
        comm_barrier_kernel<<<1, 32, 0, st>>>(p, ++c->epoch);            // all slices stored everywhere, all sends consumed
        TNK_CUDA_OK(cudaGetLastError());
        launches += 3;
    }
    if (out) TNK_CUDA_OK(cudaMemcpyAsync(out, tnk_comm_recvbuf(c), (size_t)count * 8, cudaMemcpyDeviceToDevice, st));
    if (n_launch) *n_launch = launches;
    return 0;
}
