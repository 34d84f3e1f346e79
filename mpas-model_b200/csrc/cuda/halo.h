// Halo exchange layer: replaces mpas_dmpar_exch_halo_field (src/framework/mpas_dmpar.F:5386-5552 2-D real,
// :5565-5748 3-D real) with persistent device index tables, packed device buffers and either NCCL
// grouped send/recv over NVLink (one process per GPU) or device-to-device copies (several blocks in one
// process: the reference's own no-MPI path, mpas_dmpar.F:5481-5504).
//
// Message layout per (sender, receiver): for each field of the exchange group, for each selected halo
// layer in ascending order, the elements of that layer's exchange list at position (destList-1)*inner
// (sender) / (srcList-1)*inner (receiver), exactly the reference's buffer addressing (mpas_dmpar.F:5462,
// 5521) -- i.e. ascending global ID within a layer (mpas_dmpar.F:1782-1784, 1861-1867).
#pragma once
#include <map>
#include <string>
#include <vector>

#include "common.cuh"

namespace mpasb200 {

struct HaloList {  // one reference exchange list (mpas_dmpar_types.inc mpas_exchange_list)
    int endPoint, layer;
    std::vector<int> src, dest;  // 1-based, reference semantics (see decomp.h)
};

struct HaloItem {  // one field of an exchange group
    real* array;
    int loc;    // 1 cell, 2 edge, 3 vertex
    int inner;  // product of the dimensions in front of the decomposed one
    int mask;   // bit (l-1) set = exchange halo layer l; 0 = all layers
};

struct HaloSide {                      // send or recv side of a plan
    std::vector<int> nbr;              // neighbour block ids, ascending
    std::vector<long long> off, cnt;   // per neighbour: offset / count in the packed buffer (elements)
    std::vector<int*> d_idx;           // per item: local element index (1-based) per list element
    std::vector<long long*> d_pos;     // per item: buffer offset (elements) per list element
    std::vector<long long> n;          // per item: number of list elements
    std::vector<unsigned char*> d_seg; // per item: index into nbr[] of the neighbour each list element belongs to
    real* d_buf = nullptr;
    long long total = 0;
};

struct HaloPlan {
    std::vector<HaloItem> items;
    HaloSide send, recv;
    // false: some rank's plan of this exchange does not send to exactly the ranks it receives from (or exceeds the window):
    // agreed collectively when the plan is built after the peer-to-peer switch; such an exchange travels over NCCL on every rank
    bool p2p_ok = true;
};

struct HaloBlock {  // exchange lists of one block, by location (index 1..3)
    int blockID = 0;
    std::vector<HaloList> send[4], recv[4];
    std::map<std::string, HaloPlan*> plans;
};

// NCCL through dlopen (no link-time dependency; a single-GPU run never loads it)
struct Nccl {
    void* lib = nullptr;
    void* comm = nullptr;
    int rank = 0, nranks = 1;
    bool load(std::string& err);
    int get_unique_id(void* id128);                    // fills 128 bytes
    int init_rank(const void* id128, int nranks, int rank);
    int group_start();
    int group_end();
    int send(const void* buf, size_t count, int peer, cudaStream_t s);
    int recv(void* buf, size_t count, int peer, cudaStream_t s);
    int allreduce_max(const void* in, void* out, size_t count, cudaStream_t s);
    void destroy();
    const char* last_error();
};

// ---- peer-to-peer exchange over NVLink (one process per GPU, all GPUs of one node) -----------------------------------
// The NCCL send/recv pair of an exchange costs ~40 us for these few-hundred-kilobyte messages: latency, not bandwidth.
// With the peers' receive windows mapped into this process (CUDA IPC) the pack kernel stores straight into the
// neighbour's window, one thread then publishes a sequence number per neighbour (st.release.sys) and the neighbour's
// unpack kernel waits for it (ld.acquire.sys): plain kernels, graph-capturable, no proxy thread.  The block of the pack
// kernel that finishes last publishes the numbers itself, so an exchange is two launches: pack + signal, wait + unpack.
// Window of a rank: [nranks flags (u64) padded to 256 B][nranks sources x 2 parities x cap reals].  Message m of the
// pair (src -> dst) goes to parity m & 1: src cannot send message m + 2 before it has unpacked dst's message m + 1,
// which dst sends only after unpacking m, so two slots per pair suffice.  The per-pair message counters live in device
// memory (seq_send / seq_recv), so a replayed graph advances them by itself.
#define MPAS_P2P_MAXN 32
struct P2PSideTab {  // kernel argument, one per side of a plan
    real* slot[MPAS_P2P_MAXN];                 // parity-0 slot of this pair: in the peer's window (send) / in mine (recv)
    long long off[MPAS_P2P_MAXN];              // first element of the neighbour's segment in the packed numbering
    unsigned long long* flag[MPAS_P2P_MAXN];   // send: the peer's flag for me; recv: my flag for the source
    int rank[MPAS_P2P_MAXN];                   // neighbour rank
    int n;
    long long cap;
};
struct P2P {
    bool on = false, tried = false;
    int rank = 0, nranks = 1;
    long long cap = 0;            // slot capacity (reals), the same on every rank
    char* win = nullptr;          // this rank's window
    std::vector<char*> peer;      // every rank's window as mapped here (peer[rank] == win)
    unsigned long long *seq_send = nullptr, *seq_recv = nullptr;  // device, nranks each
    int* h_timeout = nullptr;     // pinned host word, mapped: raised by a wait that gave up (that block then skipped its scatter)
    int* d_timeout = nullptr;     // the device alias of h_timeout
    int* d_timeout_dev = nullptr; // device-memory copy the spin loops poll (a host-mapped word would be a PCIe read per poll)
    unsigned* ticket = nullptr;   // device: block ticket of the pack kernel (the last block publishes the message numbers)
    unsigned long long timeout_ns = 60000000000ULL;  // MPAS_B200_TUNE_p2p_timeout_s (wall time, %globaltimer)
    static size_t flag_bytes(int nranks) { return ((size_t)nranks * 8 + 255) & ~(size_t)255; }
    real* slot(char* w, int src, int par) const { return (real*)(w + flag_bytes(nranks)) + ((size_t)src * 2 + par) * (size_t)cap; }
    unsigned long long* flag(char* w, int src) const { return (unsigned long long*)w + src; }
};
int halo_p2p_tab(const P2P& pp, const HaloSide& side, bool is_send, P2PSideTab& tab);  // non-zero: a message exceeds cap
int halo_pack_p2p(HaloPlan* p, const P2PSideTab& snd, const P2PSideTab& rcv, const P2P& pp, bool& signalled);
int halo_signal_p2p(const P2PSideTab& snd, const P2PSideTab& rcv, const P2P& pp);
int halo_unpack_p2p(HaloPlan* p, const P2PSideTab& rcv, const P2P& pp);

HaloPlan* halo_build_plan(HaloBlock& blk, const std::vector<HaloItem>& items);
void halo_free_plan(HaloPlan* p);
int halo_pack(HaloPlan* p);    // launches the pack kernels on stream()
int halo_unpack(HaloPlan* p);  // launches the unpack kernels on stream()

}  // namespace mpasb200
