// Halo exchange layer (see halo.h): plan construction from the reference-order exchange lists,
// pack / unpack kernels, NCCL binding through dlopen.
#include "halo.h"

#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <set>

namespace mpasb200 {

namespace {

// buf[pos[e] + j] = array[(idx[e]-1)*inner + j]   (PACK)   or the reverse (UNPACK); j fastest
template <bool PACK>
__global__ void __launch_bounds__(256) halo_copy_kernel(real* __restrict__ array, real* __restrict__ buf, const int* __restrict__ idx,
                                                        const long long* __restrict__ pos, long long n, int inner) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * inner) return;
    const long long e = t / inner;
    const int j = (int)(t - e * inner);
    const size_t ia = (size_t)(idx[e] - 1) * inner + j;
    const size_t ib = (size_t)pos[e] + j;
    if (PACK) buf[ib] = array[ia];
    else array[ia] = buf[ib];
}

// ---- peer-to-peer variants (halo.h): the packed buffer is the receiver's window slot of this (source, destination) pair
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// every field of the exchange group in one launch: item i covers the flat threads [start[i], start[i+1])
#define MPAS_P2P_MAXITEMS 12
struct P2PItems {
    real* array[MPAS_P2P_MAXITEMS];
    const int* idx[MPAS_P2P_MAXITEMS];
    const long long* pos[MPAS_P2P_MAXITEMS];
    const unsigned char* seg[MPAS_P2P_MAXITEMS];
    long long start[MPAS_P2P_MAXITEMS + 1];
    int inner[MPAS_P2P_MAXITEMS];
    int n;
};
__device__ __forceinline__ int p2p_find_item(const P2PItems& it, long long t) {
    int i = 0;
#pragma unroll 1
    while (i + 1 < it.n && t >= it.start[i + 1]) ++i;
    return i;
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// one thread per neighbour: count the message and publish its number to the receiver; count the messages about to be received
__device__ __forceinline__ void p2p_signal(int t, const P2PSideTab& snd, const P2PSideTab& rcv, unsigned long long* seq_send, unsigned long long* seq_recv) {
    if (t < snd.n) {
        const unsigned long long m = seq_send[snd.rank[t]] + 1ULL;
        seq_send[snd.rank[t]] = m;
        __threadfence_system();
        st_release_sys(snd.flag[t], m);
    } else if (t >= MPAS_P2P_MAXN && t - MPAS_P2P_MAXN < rcv.n) {
        seq_recv[rcv.rank[t - MPAS_P2P_MAXN]] += 1ULL;
    }
}

// window[dst][me][parity][pos - off + j] = array[(idx-1)*inner + j]; parity = (messages sent to dst so far + 1) & 1.
// SIGNAL: the block that finishes last (a ticket counter, reset for the next exchange) publishes the message numbers, so an
// exchange is two graph nodes (pack + signal, wait + unpack) instead of three.
template <bool SIGNAL>
__global__ void __launch_bounds__(256) halo_pack_p2p_kernel(P2PItems it, P2PSideTab tab, P2PSideTab rcv, unsigned long long* seq_send, unsigned long long* seq_recv,
                                                            unsigned* ticket) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < it.start[it.n]) {
        const int i = p2p_find_item(it, t);
        const long long u = t - it.start[i];
        const int inner = it.inner[i];
        const long long e = u / inner;
        const int j = (int)(u - e * inner);
        const int s = it.seg[i][e];
        const unsigned long long m = seq_send[tab.rank[s]] + 1ULL;
        real* dst = tab.slot[s] + (size_t)(m & 1ULL) * (size_t)tab.cap;
        dst[(size_t)(it.pos[i][e] - tab.off[s]) + j] = it.array[i][(size_t)(it.idx[i][e] - 1) * inner + j];
    }
    // One system-scope fence per BLOCK, not per thread: a fence.sys waits for the thread's peer stores to be acknowledged across
    // NVLink, and hundreds of thousands of them made a 5 MB exchange cost 1 ms (NCCL: 40 us).  The block barrier orders every
    // thread's stores before thread 0 (CTA scope), whose fence is cumulative: they are performed system-wide before its ticket.
    if (SIGNAL) {
        __shared__ unsigned s_last;
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();
            const unsigned k = atomicAdd(ticket, 1u);
            s_last = (k == gridDim.x - 1) ? 1u : 0u;
            if (s_last) *ticket = 0u;  // every other block has taken its ticket: ready for the next exchange
        }
        __syncthreads();
        if (s_last) {
            __threadfence();  // acquire side of the ticket: all blocks' seq_send reads and peer stores precede the signal
            p2p_signal((int)threadIdx.x, tab, rcv, seq_send, seq_recv);
        }
    }
}

__global__ void halo_signal_p2p_kernel(P2PSideTab snd, P2PSideTab rcv, unsigned long long* seq_send, unsigned long long* seq_recv) {
    p2p_signal((int)threadIdx.x, snd, rcv, seq_send, seq_recv);
}

// waits until every source of the plan has published the message expected now, then
// array[(idx-1)*inner + j] = window[me][src][parity][pos - off + j].
// A wait that gives up (timeout_ns of wall time: a peer died, or the ranks disagree about the exchange sequence) raises
// *timeout_flag -- host-mapped memory the library checks wherever it synchronises -- and the block does NOT scatter the
// stale window contents.
__global__ void __launch_bounds__(256) halo_unpack_p2p_kernel(P2PItems it, P2PSideTab tab, const unsigned long long* __restrict__ seq_recv, volatile int* timeout_flag,
                                                              volatile int* timeout_dev, unsigned long long timeout_ns) {
    __shared__ int s_fail;
    if (threadIdx.x == 0) s_fail = 0;
    __syncthreads();
    if ((int)threadIdx.x < tab.n) {
        const unsigned long long want = seq_recv[tab.rank[threadIdx.x]];
        if (ld_acquire_sys(tab.flag[threadIdx.x]) < want) {
            const unsigned long long t0 = global_timer_ns();
            while (ld_acquire_sys(tab.flag[threadIdx.x]) < want) {
                // the spin loop polls the DEVICE copy of the flag: the host-mapped word is a PCIe read, and thousands of spinning
                // threads polling it turned a 20 us exchange into 300 us
                if (*timeout_dev || global_timer_ns() - t0 > timeout_ns) {
                    *timeout_dev = 1;
                    *timeout_flag = 1;
                    __threadfence_system();
                    s_fail = 1;
                    break;
                }
                __nanosleep(100);
            }
        }
    }
    __syncthreads();
    if (s_fail) return;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= it.start[it.n]) return;
    const int i = p2p_find_item(it, t);
    const long long u = t - it.start[i];
    const int inner = it.inner[i];
    const long long e = u / inner;
    const int j = (int)(u - e * inner);
    const int s = it.seg[i][e];
    const unsigned long long m = seq_recv[tab.rank[s]];
    const real* src = tab.slot[s] + (size_t)(m & 1ULL) * (size_t)tab.cap;
    it.array[i][(size_t)(it.idx[i][e] - 1) * inner + j] = __ldcg(src + (size_t)(it.pos[i][e] - tab.off[s]) + j);  // written by the peer: not through L1
}

bool layer_selected(int mask, int layer) { return mask == 0 || (mask & (1 << (layer - 1))); }

void build_side(const std::vector<HaloItem>& items, const std::vector<HaloList> lists[4], bool is_send, HaloSide& side) {
    std::set<int> nb;
    for (const HaloItem& it : items)
        for (const HaloList& l : lists[it.loc])
            if (layer_selected(it.mask, l.layer) && !l.src.empty()) nb.insert(l.endPoint);
    side.nbr.assign(nb.begin(), nb.end());
    const size_t nI = items.size();
    std::vector<std::vector<int>> idx(nI);
    std::vector<std::vector<long long>> pos(nI);
    std::vector<std::vector<unsigned char>> seg(nI);
    long long cursor = 0;
    int inbr = -1;
    for (int n : side.nbr) {
        ++inbr;
        side.off.push_back(cursor);
        for (size_t i = 0; i < nI; ++i) {
            const HaloItem& it = items[i];
            // layers ascending; lists are stored sorted by (endPoint, layer) but do not rely on it
            std::vector<const HaloList*> sel;
            for (const HaloList& l : lists[it.loc])
                if (l.endPoint == n && layer_selected(it.mask, l.layer)) sel.push_back(&l);
            std::sort(sel.begin(), sel.end(), [](const HaloList* a, const HaloList* b) { return a->layer < b->layer; });
            for (const HaloList* l : sel) {
                for (size_t e = 0; e < l->src.size(); ++e) {
                    // send: array(:, srcList) -> buffer slot destList ; recv: buffer slot srcList -> array(:, destList)
                    const int local = is_send ? l->src[e] : l->dest[e];
                    const int slot = is_send ? l->dest[e] : l->src[e];
                    idx[i].push_back(local);
                    pos[i].push_back(cursor + (long long)(slot - 1) * it.inner);
                    seg[i].push_back((unsigned char)std::min(inbr, 255));
                }
                cursor += (long long)l->src.size() * it.inner;
            }
        }
        side.cnt.push_back(cursor - side.off.back());
    }
    side.total = cursor;
    for (size_t i = 0; i < nI; ++i) {
        int* di = nullptr;
        long long* dp = nullptr;
        unsigned char* ds = nullptr;
        side.n.push_back((long long)idx[i].size());
        if (!idx[i].empty()) {
            cudaMalloc(&di, idx[i].size() * sizeof(int));
            cudaMalloc(&dp, pos[i].size() * sizeof(long long));
            cudaMalloc(&ds, seg[i].size());
            cudaMemcpy(di, idx[i].data(), idx[i].size() * sizeof(int), cudaMemcpyHostToDevice);
            cudaMemcpy(dp, pos[i].data(), pos[i].size() * sizeof(long long), cudaMemcpyHostToDevice);
            cudaMemcpy(ds, seg[i].data(), seg[i].size(), cudaMemcpyHostToDevice);
        }
        side.d_idx.push_back(di);
        side.d_pos.push_back(dp);
        side.d_seg.push_back(ds);
    }
    if (side.total > 0) cudaMalloc(&side.d_buf, side.total * sizeof(real));
}

void free_side(HaloSide& s) {
    for (int* p : s.d_idx) cudaFree(p);
    for (long long* p : s.d_pos) cudaFree(p);
    for (unsigned char* p : s.d_seg) cudaFree(p);
    cudaFree(s.d_buf);
}

}  // namespace

HaloPlan* halo_build_plan(HaloBlock& blk, const std::vector<HaloItem>& items) {
    HaloPlan* p = new HaloPlan();
    p->items = items;
    build_side(items, blk.send, true, p->send);
    build_side(items, blk.recv, false, p->recv);
    return p;
}

void halo_free_plan(HaloPlan* p) {
    if (!p) return;
    free_side(p->send);
    free_side(p->recv);
    delete p;
}

int halo_pack(HaloPlan* p) {
    for (size_t i = 0; i < p->items.size(); ++i) {
        const long long n = p->send.n[i];
        if (n == 0) continue;
        const HaloItem& it = p->items[i];
        halo_copy_kernel<true><<<cdiv(n * it.inner, 256), 256, 0, stream()>>>(it.array, p->send.d_buf, p->send.d_idx[i], p->send.d_pos[i], n, it.inner);
        count_launch();
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail("halo_pack", e);
    return 0;
}

int halo_unpack(HaloPlan* p) {
    for (size_t i = 0; i < p->items.size(); ++i) {
        const long long n = p->recv.n[i];
        if (n == 0) continue;
        const HaloItem& it = p->items[i];
        halo_copy_kernel<false><<<cdiv(n * it.inner, 256), 256, 0, stream()>>>(it.array, p->recv.d_buf, p->recv.d_idx[i], p->recv.d_pos[i], n, it.inner);
        count_launch();
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail("halo_unpack", e);
    return 0;
}

int halo_p2p_tab(const P2P& pp, const HaloSide& side, bool is_send, P2PSideTab& tab) {
    tab.n = (int)side.nbr.size();
    tab.cap = pp.cap;
    if (tab.n > MPAS_P2P_MAXN) return 1;
    for (int s = 0; s < tab.n; ++s) {
        const int r = side.nbr[s];
        if (side.cnt[s] > pp.cap || r < 0 || r >= pp.nranks) return 1;
        tab.rank[s] = r;
        tab.off[s] = side.off[s];
        // send: slot (source = me) and flag (for me) in the destination's window; recv: slot and flag of the source in mine
        tab.slot[s] = is_send ? pp.slot(pp.peer[r], pp.rank, 0) : pp.slot(pp.win, r, 0);
        tab.flag[s] = is_send ? pp.flag(pp.peer[r], pp.rank) : pp.flag(pp.win, r);
    }
    return 0;
}

namespace {
// item descriptors of one side, in chunks of MPAS_P2P_MAXITEMS (one launch per chunk; exchange groups have <= 8 fields)
template <class F>
int p2p_for_chunks(HaloPlan* p, const HaloSide& side, F&& launch) {
    size_t i = 0;
    while (i < p->items.size()) {
        P2PItems it;
        it.n = 0;
        it.start[0] = 0;
        for (; i < p->items.size() && it.n < MPAS_P2P_MAXITEMS; ++i) {
            if (side.n[i] == 0) continue;
            const int k = it.n++;
            it.array[k] = p->items[i].array; it.idx[k] = side.d_idx[i]; it.pos[k] = side.d_pos[i]; it.seg[k] = side.d_seg[i];
            it.inner[k] = p->items[i].inner;
            it.start[k + 1] = it.start[k] + side.n[i] * p->items[i].inner;
        }
        if (it.n > 0) launch(it);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail("halo p2p launch", e);
    return 0;
}
}  // namespace

// pack (+ signal when the group is one launch: always, exchange groups have <= 8 fields); returns through `signalled`
// whether the message numbers have been published by the pack launch itself
int halo_pack_p2p(HaloPlan* p, const P2PSideTab& snd, const P2PSideTab& rcv, const P2P& pp, bool& signalled) {
    int nsel = 0;
    for (size_t i = 0; i < p->items.size(); ++i) nsel += p->send.n[i] != 0;
    signalled = pp.ticket != nullptr && nsel > 0 && nsel <= MPAS_P2P_MAXITEMS;
    const bool fold = signalled;
    return p2p_for_chunks(p, p->send, [&](const P2PItems& it) {
        if (fold) halo_pack_p2p_kernel<true><<<cdiv(it.start[it.n], 256), 256, 0, stream()>>>(it, snd, rcv, pp.seq_send, pp.seq_recv, pp.ticket);
        else halo_pack_p2p_kernel<false><<<cdiv(it.start[it.n], 256), 256, 0, stream()>>>(it, snd, rcv, pp.seq_send, pp.seq_recv, nullptr);
        count_launch();
    });
}

int halo_signal_p2p(const P2PSideTab& snd, const P2PSideTab& rcv, const P2P& pp) {
    if (snd.n == 0 && rcv.n == 0) return 0;
    halo_signal_p2p_kernel<<<1, 2 * MPAS_P2P_MAXN, 0, stream()>>>(snd, rcv, pp.seq_send, pp.seq_recv);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail("halo_signal_p2p", e);
    return 0;
}

int halo_unpack_p2p(HaloPlan* p, const P2PSideTab& rcv, const P2P& pp) {
    return p2p_for_chunks(p, p->recv, [&](const P2PItems& it) {
        halo_unpack_p2p_kernel<<<cdiv(it.start[it.n], 256), 256, 0, stream()>>>(it, rcv, pp.seq_recv, pp.d_timeout, pp.d_timeout_dev, pp.timeout_ns);
        count_launch();
    });
}

// ---- NCCL via dlopen ------------------------------------------------------------------------
namespace {
typedef struct { char internal[128]; } nccl_uid;  // ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES 128)
typedef int (*fn_getuid)(nccl_uid*);
typedef int (*fn_initrank)(void**, int, nccl_uid, int);
typedef int (*fn_void)();
typedef int (*fn_sendrecv)(void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_destroy)(void*);
typedef const char* (*fn_errstr)(int);
fn_getuid p_getuid;
fn_initrank p_initrank;
fn_void p_gstart, p_gend;
fn_sendrecv p_send, p_recv;
fn_allreduce p_allreduce;
fn_destroy p_destroy;
fn_errstr p_errstr;
int g_nccl_rc = 0;
#ifdef SINGLE_PRECISION
const int NCCL_REAL = 7;  // ncclFloat32
#else
const int NCCL_REAL = 8;  // ncclFloat64
#endif
}  // namespace

bool Nccl::load(std::string& err) {
    if (lib) return true;
    // prefer a libnccl already mapped into the process (torch's bundled one) so only one copy is live
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    if (!lib) {
        err = std::string("cannot dlopen libnccl: ") + dlerror();
        return false;
    }
    p_getuid = (fn_getuid)dlsym(lib, "ncclGetUniqueId");
    p_initrank = (fn_initrank)dlsym(lib, "ncclCommInitRank");
    p_gstart = (fn_void)dlsym(lib, "ncclGroupStart");
    p_gend = (fn_void)dlsym(lib, "ncclGroupEnd");
    p_send = (fn_sendrecv)dlsym(lib, "ncclSend");
    p_recv = (fn_sendrecv)dlsym(lib, "ncclRecv");
    p_allreduce = (fn_allreduce)dlsym(lib, "ncclAllReduce");
    p_destroy = (fn_destroy)dlsym(lib, "ncclCommDestroy");
    p_errstr = (fn_errstr)dlsym(lib, "ncclGetErrorString");
    if (!p_getuid || !p_initrank || !p_gstart || !p_gend || !p_send || !p_recv || !p_destroy) {
        err = "libnccl lacks a required symbol";
        return false;
    }
    return true;
}
int Nccl::get_unique_id(void* id128) { return g_nccl_rc = p_getuid((nccl_uid*)id128); }
int Nccl::init_rank(const void* id128, int nranks_, int rank_) {
    nccl_uid id;
    std::memcpy(&id, id128, sizeof(id));
    rank = rank_;
    nranks = nranks_;
    return g_nccl_rc = p_initrank(&comm, nranks_, id, rank_);
}
int Nccl::group_start() { return g_nccl_rc = p_gstart(); }
int Nccl::group_end() { return g_nccl_rc = p_gend(); }
int Nccl::send(const void* buf, size_t count, int peer, cudaStream_t s) { return g_nccl_rc = p_send((void*)buf, count, NCCL_REAL, peer, comm, s); }
int Nccl::recv(void* buf, size_t count, int peer, cudaStream_t s) { return g_nccl_rc = p_recv(buf, count, NCCL_REAL, peer, comm, s); }
int Nccl::allreduce_max(const void* in, void* out, size_t count, cudaStream_t s) {
    return g_nccl_rc = p_allreduce(in, out, count, NCCL_REAL, 2 /* ncclMax */, comm, s);
}
void Nccl::destroy() {
    if (comm && p_destroy) p_destroy(comm);
    comm = nullptr;
}
const char* Nccl::last_error() { return p_errstr ? p_errstr(g_nccl_rc) : "nccl error"; }

}  // namespace mpasb200
