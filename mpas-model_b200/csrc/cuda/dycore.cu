// Device-resident dycore: field store, halo-exchange groups, C ABI (mpas_b200_dycore_*).
// The step sequence itself is in srk3.cu.
#include "dycore.h"

#include <algorithm>
#include <cstring>
#include <stdexcept>

#include "../host/decomp.h"

namespace mpasb200 {
extern std::string g_last_error;  // runtime.cu
}

namespace mpasb200 {

real* DBlock::R(const std::string& n) {
    auto it = f.find(n);
    if (it == f.end()) throw std::runtime_error("dycore: missing field " + n);
    return (real*)it->second.dev;
}
real* DBlock::Ropt(const std::string& n) {
    auto it = f.find(n);
    return it == f.end() ? nullptr : (real*)it->second.dev;
}
int* DBlock::I(const std::string& n) {
    auto it = f.find(n);
    if (it == f.end()) throw std::runtime_error("dycore: missing field " + n);
    return (int*)it->second.dev;
}
double DBlock::c(const std::string& n) const {
    auto it = cfg.find(n);
    if (it == cfg.end()) throw std::runtime_error("dycore: missing config " + n);
    return it->second;
}
double DBlock::c(const std::string& n, double dflt) const {
    auto it = cfg.find(n);
    return it == cfg.end() ? dflt : it->second;
}

// One exchange of a group of fields (adjacent reference exchanges with no compute in between are
// aggregated into one message per neighbour).
int dycore_exchange(Dycore& m, const std::vector<XField>& fields) {
    m.nExchanges += (long)fields.size();
    const int P = (int)m.blocks.size();
    if (P == 1 && m.nranks == 1) return 0;
    std::string key;
    for (const XField& xf : fields) key += xf.name + ":" + std::to_string(xf.mask) + ";";
    std::vector<HaloPlan*> plans(P);
    for (int b = 0; b < P; ++b) {
        DBlock& blk = m.blocks[b];
        // the plan caches array addresses; time levels swap every step, so the key carries the address
        std::string k2 = key;
        for (const XField& xf : fields) k2 += std::to_string((uintptr_t)blk.R(xf.name)) + ",";
        auto it = blk.halo.plans.find(k2);
        if (it == blk.halo.plans.end()) {
            std::vector<HaloItem> items;
            for (const XField& xf : fields) {
                const DField& df = blk.f.at(xf.name);
                if (df.loc < 1 || df.loc > 3) throw std::runtime_error("dycore_exchange: field " + xf.name + " is not decomposed");
                HaloItem hi;
                hi.array = (real*)df.dev; hi.loc = df.loc; hi.inner = df.inner; hi.mask = xf.mask;
                items.push_back(hi);
            }
            HaloPlan* p = halo_build_plan(blk.halo, items);
            blk.halo.plans[k2] = p;
            plans[b] = p;
            if (m.nranks > 1 && m.p2p.on && P == 1) {
                // a plan first seen after the switch to peer stores (every rank builds it at the same call): all ranks agree
                // whether it fits the two-slot protocol; if it does not on ANY rank it goes over NCCL on all of them
                P2PSideTab a, c;
                const bool fits = p->send.nbr == p->recv.nbr && !halo_p2p_tab(m.p2p, p->send, true, a) && !halo_p2p_tab(m.p2p, p->recv, false, c);
                real hv = fits ? (real)0.0 : (real)1.0;
                real* d_tmp = nullptr;
                if (cudaMalloc(&d_tmp, 2 * sizeof(real)) != cudaSuccess) return fail_msg("dycore_exchange: cudaMalloc");
                cudaMemcpyAsync(d_tmp, &hv, sizeof(real), cudaMemcpyHostToDevice, stream());
                const int nrc = m.nccl.allreduce_max(d_tmp, d_tmp + 1, 1, stream());
                cudaMemcpyAsync(&hv, d_tmp + 1, sizeof(real), cudaMemcpyDeviceToHost, stream());
                const cudaError_t se = cudaStreamSynchronize(stream());
                cudaFree(d_tmp);
                if (nrc || se != cudaSuccess) return fail_msg("dycore_exchange: collective check of a new exchange plan failed (a plan cannot be built inside a stream capture)");
                p->p2p_ok = hv == (real)0.0;
            }
        } else {
            plans[b] = it->second;
        }
    }
    if (m.nranks > 1 && m.p2p.on && P == 1 && plans[0]->p2p_ok) {
        // peer-to-peer: pack straight into the neighbours' windows, publish, wait + unpack (halo.h)
        HaloPlan* p = plans[0];
        P2PSideTab snd, rcv;
        if (p->send.nbr != p->recv.nbr || halo_p2p_tab(m.p2p, p->send, true, snd) || halo_p2p_tab(m.p2p, p->recv, false, rcv))
            return fail_msg("dycore_exchange: exchange '" + key + "' does not fit the peer-to-peer protocol (message larger than the window, or it does not "
                            "send to exactly the ranks it receives from); run with MPAS_B200_TUNE_p2p=0");
        bool signalled = false;
        int rc = halo_pack_p2p(p, snd, rcv, m.p2p, signalled);
        if (!rc && !signalled) rc = halo_signal_p2p(snd, rcv, m.p2p);
        if (!rc) rc = halo_unpack_p2p(p, rcv, m.p2p);
        return rc;
    }
    for (int b = 0; b < P; ++b) {
        int rc = halo_pack(plans[b]);
        if (rc) return rc;
    }
    if (m.nranks > 1) {
        if (!m.have_comm) return fail_msg("dycore_exchange: nranks > 1 but no NCCL communicator (call mpas_b200_dycore_comm_init)");
        HaloPlan* p = plans[0];
        if (m.nccl.group_start()) return fail_msg(std::string("ncclGroupStart: ") + m.nccl.last_error());
        for (size_t i = 0; i < p->recv.nbr.size(); ++i)
            if (m.nccl.recv(p->recv.d_buf + p->recv.off[i], (size_t)p->recv.cnt[i], p->recv.nbr[i], stream())) return fail_msg(std::string("ncclRecv: ") + m.nccl.last_error());
        for (size_t i = 0; i < p->send.nbr.size(); ++i)
            if (m.nccl.send(p->send.d_buf + p->send.off[i], (size_t)p->send.cnt[i], p->send.nbr[i], stream())) return fail_msg(std::string("ncclSend: ") + m.nccl.last_error());
        if (m.nccl.group_end()) return fail_msg(std::string("ncclGroupEnd: ") + m.nccl.last_error());
        count_launch();
    } else {
        // blocks of one process: message = device copy from the sender's packed segment to the receiver's
        for (int b = 0; b < P; ++b) {
            HaloPlan* ps = plans[b];
            for (size_t i = 0; i < ps->send.nbr.size(); ++i) {
                const int dstBlock = ps->send.nbr[i];
                if (dstBlock < 0 || dstBlock >= P) return fail_msg("dycore_exchange: neighbour block not in this process");
                HaloPlan* pr = plans[dstBlock];
                size_t j = 0;
                for (; j < pr->recv.nbr.size(); ++j)
                    if (pr->recv.nbr[j] == m.blocks[b].halo.blockID) break;
                if (j == pr->recv.nbr.size() || pr->recv.cnt[j] != ps->send.cnt[i]) return fail_msg("dycore_exchange: unmatched exchange lists");
                cudaError_t e = cudaMemcpyAsync(pr->recv.d_buf + pr->recv.off[j], ps->send.d_buf + ps->send.off[i], (size_t)ps->send.cnt[i] * sizeof(real),
                                                cudaMemcpyDeviceToDevice, stream());
                if (e != cudaSuccess) return fail("dycore_exchange copy", e);
                count_launch();
            }
        }
    }
    for (int b = 0; b < P; ++b) {
        int rc = halo_unpack(plans[b]);
        if (rc) return rc;
    }
    return 0;
}

// A peer-to-peer wait that gave up leaves the halo columns of that exchange stale: every entry point that has just
// synchronised (or is about to start new work) turns the raised flag into an error instead of carrying on.
int dycore_check_p2p(Dycore& m) {
    if (m.p2p.on && m.p2p.h_timeout && *(volatile int*)m.p2p.h_timeout)
        return fail_msg("peer-to-peer halo exchange timed out waiting for a neighbour rank (MPAS_B200_TUNE_p2p_timeout_s, default 60 s): the state of this "
                        "rank is no longer valid; a rank died or the ranks disagree about the exchange sequence");
    return 0;
}

// ---- peer-to-peer windows ------------------------------------------------------------------------------------------------
void dycore_p2p_teardown(Dycore& m) {
    P2P& pp = m.p2p;
    for (int r = 0; r < (int)pp.peer.size(); ++r)
        if (pp.peer[r] && pp.peer[r] != pp.win) cudaIpcCloseMemHandle(pp.peer[r]);
    pp.peer.clear();
    if (pp.win) cudaFree(pp.win);
    if (pp.seq_send) cudaFree(pp.seq_send);
    if (pp.seq_recv) cudaFree(pp.seq_recv);
    if (pp.h_timeout) cudaFreeHost(pp.h_timeout);
    if (pp.ticket) cudaFree(pp.ticket);
    if (pp.d_timeout_dev) cudaFree(pp.d_timeout_dev);
    pp.d_timeout_dev = nullptr;
    pp.win = nullptr; pp.seq_send = pp.seq_recv = nullptr; pp.d_timeout = pp.h_timeout = nullptr; pp.ticket = nullptr;
    pp.on = false;
}

int dycore_p2p_setup(Dycore& m) {
    P2P& pp = m.p2p;
    if (pp.tried) return 0;
    pp.tried = true;
    if (m.nranks < 2 || !m.have_comm || m.blocks.size() != 1 || m.nranks > MPAS_P2P_MAXN || !tune("p2p", 1)) return 0;
    pp.rank = m.rank;
    pp.nranks = m.nranks;
    // slot capacity: the largest message of any plan built so far (both time-level parities have run), on any rank
    long long cap = 0;
    for (auto& kv : m.blocks[0].halo.plans) {
        for (long long c : kv.second->send.cnt) cap = std::max(cap, c);
        for (long long c : kv.second->recv.cnt) cap = std::max(cap, c);
    }
    cap = (cap + 65535) & ~65535LL;  // a multiple of 2^16 is exact in the float of the allreduce of a single-precision build
    // the two-slot protocol relies on pairwise back-pressure: every plan must send to exactly the ranks it receives from
    bool symmetric = true;
    for (auto& kv : m.blocks[0].halo.plans) symmetric = symmetric && kv.second->send.nbr == kv.second->recv.nbr;
    const int HW = 64 / (int)sizeof(real);  // reals per IPC handle
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    real* d_tmp = nullptr;
    if (cudaMalloc(&d_tmp, (size_t)(2 + (m.nranks + 1) * HW) * sizeof(real)) != cudaSuccess) return fail_msg("dycore_p2p_setup: cudaMalloc");
    auto allmax = [&](double v, double& out) -> int {
        real hv = (real)v;
        cudaMemcpyAsync(d_tmp, &hv, sizeof(real), cudaMemcpyHostToDevice, stream());
        if (m.nccl.allreduce_max(d_tmp, d_tmp + 1, 1, stream())) return 1;
        cudaMemcpyAsync(&hv, d_tmp + 1, sizeof(real), cudaMemcpyDeviceToHost, stream());
        if (cudaStreamSynchronize(stream()) != cudaSuccess) return 1;
        out = (double)hv;
        return 0;
    };
    double gcap = 0;
    if (allmax((double)cap, gcap)) { cudaFree(d_tmp); return fail_msg(std::string("dycore_p2p_setup: allreduce: ") + m.nccl.last_error()); }
    pp.cap = (long long)gcap;
    bool ok = pp.cap > 0 && symmetric;
    // window + counters
    const size_t wbytes = P2P::flag_bytes(m.nranks) + (size_t)m.nranks * 2 * (size_t)pp.cap * sizeof(real);
    cudaIpcMemHandle_t mine;
    std::memset(&mine, 0, sizeof(mine));
    if (ok) ok = cudaMalloc(&pp.win, wbytes) == cudaSuccess;
    if (ok) ok = cudaMemsetAsync(pp.win, 0, wbytes, stream()) == cudaSuccess;
    if (ok) ok = cudaIpcGetMemHandle(&mine, pp.win) == cudaSuccess;
    if (ok) ok = cudaMalloc(&pp.seq_send, m.nranks * sizeof(unsigned long long)) == cudaSuccess && cudaMalloc(&pp.seq_recv, m.nranks * sizeof(unsigned long long)) == cudaSuccess &&
                 cudaMalloc(&pp.ticket, sizeof(unsigned)) == cudaSuccess && cudaMalloc(&pp.d_timeout_dev, sizeof(int)) == cudaSuccess &&
                 cudaHostAlloc(&pp.h_timeout, sizeof(int), cudaHostAllocMapped) == cudaSuccess;
    if (ok) {
        *pp.h_timeout = 0;
        ok = cudaHostGetDevicePointer(&pp.d_timeout, pp.h_timeout, 0) == cudaSuccess;
        pp.timeout_ns = (unsigned long long)std::max(1, tune("p2p_timeout_s", 60)) * 1000000000ULL;
        if (!tune("p2p_fold_signal", 1)) { cudaFree(pp.ticket); pp.ticket = nullptr; }
    }
    if (ok) {
        cudaMemsetAsync(pp.seq_send, 0, m.nranks * sizeof(unsigned long long), stream());
        cudaMemsetAsync(pp.seq_recv, 0, m.nranks * sizeof(unsigned long long), stream());
        if (pp.ticket) cudaMemsetAsync(pp.ticket, 0, sizeof(unsigned), stream());
        cudaMemsetAsync(pp.d_timeout_dev, 0, sizeof(int), stream());
    }
    cudaGetLastError();
    // every rank's handle to every rank (a failed rank still takes part, with a zero handle)
    real* d_mine = d_tmp + 2;
    real* d_all = d_tmp + 2 + HW;
    cudaMemcpyAsync(d_mine, &mine, 64, cudaMemcpyHostToDevice, stream());
    int nrc = m.nccl.group_start();
    for (int r = 0; r < m.nranks && !nrc; ++r) {
        if (r == m.rank) continue;
        nrc = m.nccl.recv(d_all + (size_t)r * HW, HW, r, stream());
        if (!nrc) nrc = m.nccl.send(d_mine, HW, r, stream());
    }
    if (!nrc) nrc = m.nccl.group_end();
    std::vector<cudaIpcMemHandle_t> all(m.nranks);
    cudaMemcpyAsync(all.data(), d_all, (size_t)m.nranks * 64, cudaMemcpyDeviceToHost, stream());
    if (cudaStreamSynchronize(stream()) != cudaSuccess || nrc) ok = false;
    pp.peer.assign(m.nranks, nullptr);
    if (ok) {
        pp.peer[m.rank] = pp.win;
        for (int r = 0; r < m.nranks && ok; ++r) {
            if (r == m.rank) continue;
            void* q = nullptr;
            ok = cudaIpcOpenMemHandle(&q, all[r], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            pp.peer[r] = (char*)q;
        }
    }
    cudaGetLastError();
    // all or nothing (also the barrier that orders every rank's window clear before the first peer store)
    double anyfail = 1;
    const int arc = allmax(ok ? 0.0 : 1.0, anyfail);
    cudaFree(d_tmp);
    if (arc || anyfail != 0.0) {
        dycore_p2p_teardown(m);
        return 0;  // stay on NCCL
    }
    pp.on = true;
    return 0;
}

namespace {
struct ZeroTab {
    real* p[16];
    int n[16];
};
__global__ void zero_cols_kernel(ZeroTab t) {
    real* p = t.p[blockIdx.x];
    for (int i = threadIdx.x; i < t.n[blockIdx.x]; i += blockDim.x) p[i] = (real)0.0;
}
}  // namespace

// The reference zero-fills the garbage column of its scratch arrays every time it allocates them (TI:593-605, 1029-1050);
// views of the shared arena get the same at the start of their phase.  group 2 also covers the all-scalars workspace (4).
int dycore_zero_garbage(DBlock& b, int group) {
    if (b.views.empty()) return 0;
    ZeroTab t;
    int k = 0;
    for (auto& v : b.views) {
        if (v.group != group && !(group == 2 && v.group == 4)) continue;
        t.p[k] = v.garbage; t.n[k] = v.inner;
        if (++k == 16) {
            zero_cols_kernel<<<k, 128, 0, stream()>>>(t);
            count_launch();
            k = 0;
        }
    }
    if (k) {
        zero_cols_kernel<<<k, 128, 0, stream()>>>(t);
        count_launch();
    }
    return 0;
}

namespace {

int add_field(DBlock& blk, const std::string& name, const void* host, int kind, int loc, const std::vector<long long>& dims) {
    DField df;
    df.kind = kind; df.loc = loc; df.dims = dims;
    size_t n = 1;
    for (auto d : dims) n *= (size_t)d;
    df.inner = 1;
    for (size_t k = 0; k + 1 < dims.size(); ++k) df.inner *= (int)dims[k];
    df.bytes = n * (kind == 0 ? sizeof(real) : sizeof(int));
    cudaError_t e = cudaMalloc(&df.dev, df.bytes + 64);  // +64: bulk column copies read a 16-byte aligned window around a column
    if (e != cudaSuccess) return fail(("cudaMalloc " + name).c_str(), e);
    if (host) {
        if (kind == 0 && sizeof(real) != sizeof(double)) {
            std::vector<real> tmp(n);
            const double* src = (const double*)host;
            for (size_t i = 0; i < n; ++i) tmp[i] = (real)src[i];
            e = cudaMemcpy(df.dev, tmp.data(), df.bytes, cudaMemcpyHostToDevice);
        } else {
            e = cudaMemcpy(df.dev, host, df.bytes, cudaMemcpyHostToDevice);
        }
        if (e != cudaSuccess) return fail(("upload " + name).c_str(), e);
    } else {
        cudaMemset(df.dev, 0, df.bytes);
    }
    auto it = blk.f.find(name);
    if (it != blk.f.end() && it->second.owned) cudaFree(it->second.dev);
    blk.f[name] = df;
    return 0;
}

void copy_lists(const std::vector<mpas::ExchangeList>& in, std::vector<HaloList>& out) {
    for (const auto& l : in) {
        HaloList h;
        h.endPoint = l.endPointID; h.layer = l.haloLayer; h.src = l.srcList; h.dest = l.destList;
        out.push_back(h);
    }
}

}  // namespace
}  // namespace mpasb200

using namespace mpasb200;

#define DY_TRY try {
#define DY_CATCH                                 \
    }                                            \
    catch (const std::exception& e) {            \
        mpasb200::g_last_error = e.what();           \
        return 1;                                \
    }

extern "C" {

void* mpas_b200_dycore_create(int nLocalBlocks, int rank, int nranks) {
    Dycore* m = new Dycore();
    m->blocks.resize(nLocalBlocks);
    m->rank = rank;
    m->nranks = nranks;
    stream();
    return m;
}

void mpas_b200_dycore_destroy(void* h) {
    Dycore* m = (Dycore*)h;
    if (!m) return;
    cudaStreamSynchronize(stream());
    dycore_p2p_teardown(*m);
    for (auto& b : m->blocks) {
        for (const char* key : {"scale_arr", "tend_u"}) {  // the library workspace keyed by this block's arrays (runtime.cu)
            auto sa = b.f.find(key);
            if (sa != b.f.end()) release_keyed_workspace(sa->second.dev);
        }
        if (b.arena) cudaFree(b.arena);
        auto adv = b.f.find("advCellsForEdge");
        if (adv != b.f.end()) release_edge_tile_plans(adv->second.dev);  // cached by the device address of this array (tiles.cu)
        for (auto& kv : b.f)
            if (kv.second.owned) cudaFree(kv.second.dev);
        for (auto& kv : b.halo.plans) halo_free_plan(kv.second);
    }
    for (int i = 0; i < 2; ++i)
        if (m->graph_exec[i]) cudaGraphExecDestroy(m->graph_exec[i]);
    for (auto e : m->event_pool) cudaEventDestroy(e);
    if (m->d_reduce) cudaFree(m->d_reduce);
    if (m->d_minmax) cudaFree(m->d_minmax);
    m->nccl.destroy();
    delete m;
}

int mpas_b200_dycore_attach_block(void* h, int iLocal, void* hostBlock) {
    DY_TRY
    Dycore* m = (Dycore*)h;
    mpas::Block* hb = (mpas::Block*)hostBlock;
    if (!hb) return fail_msg("mpas_b200_dycore_attach_block: null host block");
    DBlock& blk = m->blocks.at(iLocal);
    // diet: (a) arrays only physics or output use -- every kernel takes them as optional arguments -- get no mirror;
    // (b) phase-local scratch becomes views of one arena sized for the largest phase
    static const char* kOptional[] = {"gradPVn", "gradPVt", "ke_edge", "tend_ru_physics", "rthdynten", "tend_rtheta_adv", "tend_rho_physics",
                                      "tend_rtheta_physics", "rt_diabatic_tend"};
    static const std::map<std::string, int> kGroup = {
        {"delsq_u", 1}, {"delsq_theta", 1}, {"delsq_w", 1}, {"delsq_divergence", 1}, {"dpdz", 1}, {"delsq_vorticity", 1},
        {"scalar_old", 2}, {"scalar_new", 2}, {"s_max", 2}, {"s_min", 2}, {"scale_arr", 2}, {"wdtn", 2}, {"flux_arr", 2}, {"flux_upwind_tmp", 2}, {"flux_tmp", 2},
        {"horiz_flux_arr", 3}};
    struct Pending { std::string name; int group; size_t off; DField df; };
    std::vector<Pending> pend;
    size_t gsize[4] = {0, 0, 0, 0};
    auto place = [&](int g, size_t bytes) {  // 256-byte aligned, 64 bytes of slack behind (16-byte windows of the bulk column copies)
        const size_t off = gsize[g];
        gsize[g] = (off + bytes + 64 + 255) & ~(size_t)255;
        return off;
    };
    for (auto& kv : hb->pool.f) {
        const mpas::Field& fl = kv.second;
        std::vector<long long> dims(fl.dims.begin(), fl.dims.end());
        if (m->diet && fl.kind == mpas::FK_REAL) {
            bool skip = false;
            for (const char* o : kOptional) skip = skip || kv.first == o;
            if (skip) continue;
            auto g = kGroup.find(kv.first);
            if (g != kGroup.end()) {
                Pending pd;
                pd.name = kv.first; pd.group = g->second;
                pd.df.kind = 0; pd.df.loc = (int)fl.loc; pd.df.dims = dims; pd.df.owned = false;
                size_t n = 1;
                for (auto d : dims) n *= (size_t)d;
                pd.df.inner = 1;
                for (size_t k = 0; k + 1 < dims.size(); ++k) pd.df.inner *= (int)dims[k];
                pd.df.bytes = n * sizeof(real);
                pd.off = place(pd.group, pd.df.bytes);
                pend.push_back(pd);
                continue;
            }
        }
        const void* src = fl.kind == mpas::FK_REAL ? (const void*)fl.r.data() : (const void*)fl.i.data();
        if (fl.kind == mpas::FK_REAL && fl.zero && fl.r.empty()) src = nullptr;  // lean block: zero-filled mirror from the metadata alone
        int rc = add_field(blk, kv.first, src, fl.kind == mpas::FK_REAL ? 0 : 1, (int)fl.loc, dims);
        if (rc) return rc;
    }
    for (auto& kv : hb->pool.cfg) blk.cfg[kv.first] = kv.second;
    if (m->diet && !pend.empty()) {
        const int L = (int)blk.c("nVertLevels"), ns = (int)blk.c("num_scalars"), nC = (int)blk.c("nCells"), nE = (int)blk.c("nEdges");
        // library workspace of the same phases, as views too: the three per-edge flux arrays of dyn_tend (group 1, behind the rk 1
        // scratch) and the all-scalars mono workspace (group 2, OVER the per-scalar scratch: a call takes one path or the other)
        const size_t nELp = ((size_t)(nE + 1) * L + 1) & ~(size_t)1;
        const size_t off_flux = place(1, 3 * nELp * sizeof(real));
        const MonoWsLayout lay(L, ns, nC, nE);
        gsize[2] = std::max(gsize[2], (lay.total * sizeof(real) + 64 + 255) & ~(size_t)255);
        blk.arena_bytes = std::max(gsize[1], std::max(gsize[2], gsize[3]));
        cudaError_t e = cudaMalloc(&blk.arena, blk.arena_bytes);
        if (e != cudaSuccess) return fail("cudaMalloc scratch arena", e);
        cudaMemset(blk.arena, 0, blk.arena_bytes);
        for (auto& pd : pend) {
            pd.df.dev = blk.arena + pd.off;
            blk.f[pd.name] = pd.df;
            ArenaView v;
            v.name = pd.name; v.group = pd.group; v.inner = pd.df.inner;
            v.garbage = (real*)pd.df.dev + (size_t)pd.df.inner * (size_t)(pd.df.dims.back() - 1);
            blk.views.push_back(v);
        }
        // garbage columns of the all-scalars workspace pieces (group 2)
        {
            real* ws = (real*)blk.arena;
            const int nel = L * ns, nel1 = (L + 1) * ns;
            real* base[5] = {ws, ws + lay.n_edge, ws + 2 * lay.n_edge, ws + 2 * lay.n_edge + lay.n_wd, ws + 2 * lay.n_edge + lay.n_wd + lay.n_cell};
            const int inner[5] = {nel, nel, nel1, nel, 2 * nel};
            const int last[5] = {nE, nE, nC, nC, nC};
            const char* nm[5] = {"mono_ws_flux", "mono_ws_upwind", "mono_ws_wdtn", "mono_ws_snew", "mono_ws_scale"};
            for (int i = 0; i < 5; ++i) blk.views.push_back({nm[i], 4, base[i] + (size_t)inner[i] * last[i], inner[i]});
        }
        provide_keyed_workspace(blk.f.at("tend_u").dev, 1, (real*)(blk.arena + off_flux), 3 * nELp);
        provide_keyed_workspace(blk.f.at("scale_arr").dev, 0, (real*)blk.arena, lay.total);
    }
    blk.halo.blockID = hb->blockID;
    copy_lists(hb->cellExch.send, blk.halo.send[1]); copy_lists(hb->cellExch.recv, blk.halo.recv[1]);
    copy_lists(hb->edgeExch.send, blk.halo.send[2]); copy_lists(hb->edgeExch.recv, blk.halo.recv[2]);
    copy_lists(hb->vertexExch.send, blk.halo.send[3]); copy_lists(hb->vertexExch.recv, blk.halo.recv[3]);
    return 0;
    DY_CATCH
}

// 1 when `name` of block iLocal is a view of the scratch arena (valid only inside its phase), 0 when it has its own storage, -1 when absent
int mpas_b200_dycore_field_is_scratch_view(void* h, int iLocal, const char* name) {
    DBlock& b = ((Dycore*)h)->blocks.at(iLocal);
    auto it = b.f.find(name);
    if (it == b.f.end()) return -1;
    for (auto& v : b.views)
        if (v.name == name) return 1;
    return 0;
}

long long mpas_b200_dycore_device_bytes(void* h) {
    long long n = 0;
    for (auto& b : ((Dycore*)h)->blocks) {
        n += (long long)b.arena_bytes;
        for (auto& kv : b.f)
            if (kv.second.owned) n += (long long)kv.second.bytes + 64;
    }
    return n;
}

int mpas_b200_dycore_set_cfg(void* h, int iLocal, const char* name, double value) {
    ((Dycore*)h)->blocks.at(iLocal).cfg[name] = value;
    return 0;
}
double mpas_b200_dycore_get_cfg(void* h, int iLocal, const char* name) { return ((Dycore*)h)->blocks.at(iLocal).c(name, 0.0); }

int mpas_b200_nccl_unique_id(void* id128) {
    Nccl n;
    std::string err;
    if (!n.load(err)) return fail_msg(err);
    if (n.get_unique_id(id128)) return fail_msg(std::string("ncclGetUniqueId: ") + n.last_error());
    return 0;
}

int mpas_b200_dycore_comm_init(void* h, const void* id128) {
    Dycore* m = (Dycore*)h;
    std::string err;
    if (!m->nccl.load(err)) return fail_msg(err);
    if (m->nccl.init_rank(id128, m->nranks, m->rank)) return fail_msg(std::string("ncclCommInitRank: ") + m->nccl.last_error());
    m->have_comm = true;
    return 0;
}

int mpas_b200_dycore_init(void* h) {
    DY_TRY
    return dycore_init(*(Dycore*)h);
    DY_CATCH
}

int mpas_b200_dycore_step(void* h, int nsteps) {
    DY_TRY
    Dycore& m = *(Dycore*)h;
    for (int i = 0; i < nsteps; ++i) {
        int rc = dycore_check_p2p(m);
        if (!rc) rc = dycore_step(m);
        if (rc) return rc;
    }
    return 0;
    DY_CATCH
}

int mpas_b200_dycore_get_field(void* h, int iLocal, const char* name, void* host_out, size_t bytes) {
    DY_TRY
    DBlock& b = ((Dycore*)h)->blocks.at(iLocal);
    auto it = b.f.find(name);
    if (it == b.f.end()) return fail_msg(std::string("no field ") + name);
    if (bytes != it->second.bytes) return fail_msg(std::string("size mismatch for field ") + name);
    cudaError_t e = cudaMemcpyAsync(host_out, it->second.dev, bytes, cudaMemcpyDeviceToHost, stream());
    if (e != cudaSuccess) return fail("mpas_b200_dycore_get_field", e);
    e = cudaStreamSynchronize(stream());
    if (e != cudaSuccess) return fail("mpas_b200_dycore_get_field sync", e);
    return dycore_check_p2p(*(Dycore*)h);
    DY_CATCH
}

int mpas_b200_dycore_set_field(void* h, int iLocal, const char* name, const void* host_in, size_t bytes) {
    DY_TRY
    DBlock& b = ((Dycore*)h)->blocks.at(iLocal);
    auto it = b.f.find(name);
    if (it == b.f.end()) return fail_msg(std::string("no field ") + name);
    if (bytes != it->second.bytes) return fail_msg(std::string("size mismatch for field ") + name);
    cudaError_t e = cudaMemcpyAsync(it->second.dev, host_in, bytes, cudaMemcpyHostToDevice, stream());
    if (e != cudaSuccess) return fail("mpas_b200_dycore_set_field", e);
    return 0;
    DY_CATCH
}

void* mpas_b200_dycore_field_device_ptr(void* h, int iLocal, const char* name) {
    DBlock& b = ((Dycore*)h)->blocks.at(iLocal);
    auto it = b.f.find(name);
    return it == b.f.end() ? nullptr : it->second.dev;
}

int mpas_b200_dycore_exchange(void* h, const char* name, int layer) {
    DY_TRY
    return dycore_exchange(*(Dycore*)h, {{name, layer == 0 ? 0 : (1 << (layer - 1))}});
    DY_CATCH
}

int mpas_b200_dycore_call(void* h, const char* routine, const double* params, int nparams) {
    DY_TRY
    return dycore_call(*(Dycore*)h, routine, params, nparams);
    DY_CATCH
}

long mpas_b200_dycore_exchange_count(void* h) { return ((Dycore*)h)->nExchanges; }

// 0: no inter-process exchange (one rank), 1: NCCL grouped send/recv, 2: peer-to-peer stores into mapped windows.
// Negative: a peer-to-peer wait timed out since the last call (a rank died or the ranks disagree about the sequence).
int mpas_b200_dycore_halo_transport(void* h) {
    Dycore* m = (Dycore*)h;
    if (m->nranks < 2) return 0;
    if (!m->p2p.on) return 1;
    cudaStreamSynchronize(stream());
    if (*(volatile int*)m->p2p.h_timeout) return -2;
    return 2;
}

int mpas_b200_dycore_set_option(void* h, const char* name, int value) {
    Dycore* m = (Dycore*)h;
    std::string n = name;
    if (n == "timing") m->timing = value != 0;
    else if (n == "graph") m->use_graph = value != 0;
    else if (n == "aggregate") m->aggregate = value != 0;
    else if (n == "diet") m->diet = value != 0;  // before attach_block
    else if (n == "fold_first") m->fold_first = value != 0;  // before the first step (a captured graph keeps what it was captured with)
    else return fail_msg("unknown option " + n);
    return 0;
}

int mpas_b200_dycore_timer_count(void* h) { return (int)((Dycore*)h)->timer_order.size(); }
int mpas_b200_dycore_timer_get(void* h, int idx, char* name, int cap, double* ms, long* calls) {
    Dycore* m = (Dycore*)h;
    if (idx < 0 || idx >= (int)m->timer_order.size()) return 1;
    const std::string& n = m->timer_order[idx];
    std::strncpy(name, n.c_str(), cap - 1);
    name[cap - 1] = 0;
    PhaseTimer& t = m->timers[n];
    *ms = t.ms;
    *calls = t.calls;
    return 0;
}
int mpas_b200_dycore_timer_reset(void* h) {
    Dycore* m = (Dycore*)h;
    for (auto& kv : m->timers) { kv.second.ms = 0; kv.second.calls = 0; }
    return 0;
}

}  // extern "C"
