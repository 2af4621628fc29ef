// Redistribution engine: dst := src (or src^T) and dst += alpha * sum(src partials)
// between any two supported element-cyclic distributions with arbitrary alignments.
//
// What it replaces in the reference (per (dst,src) pair one of, or a chain of):
//   copy::RowAllGather / ColAllGather / Partial{Row,Col}AllGather / {Row,Col}AllToAllDemote /
//   {Row,Col}AllToAllPromote / Exchange / {Col,Row}wiseVectorExchange / *Filter / TransposeDist
//     (include/El/blas_like/level1/Copy/*.hpp, dispatch tables src/core/DistMatrix/ElementMatrix/*.cpp),
//   dist Transpose (include/El/blas_like/level1/Transpose.hpp:191-250, Transpose/*.hpp),
//   AxpyContract / Contract / TransposeAxpyContract (AxpyContract.hpp:14-544, Contract.hpp,
//     TransposeAxpyContract.hpp:14-48).
//
// B200-first design: every such operation is ONE plan of per-peer lattices (closed-form CRT
// intersection of the source rank's and the destination rank's index progressions, dist.h),
// executed as  pack kernel (all peers, one launch)  ->  one NCCL group of send/recv over
// NVSwitch  ->  unpack kernel (all peers, one launch).  Multi-hop chains of the reference
// (e.g. [MC,MR] -> [VC,*] -> [VR,*] -> transpose -> [*,MR]) collapse to one hop because every
// peer is one NVSwitch hop away; the transpose is folded into the pack.  The element
// permutation is identical to the reference's, so results are bit-exact.
#include "grid.h"
#include <stdlib.h>
#include "hyd_internal.h"
#include "lattice.cuh"
#include <algorithm>
#include <vector>

namespace hyd {

std::atomic<long long> g_launches{0}, g_allocs{0}, g_syncs{0};

int ws_reserve(Workspace& w, size_t bytes) {
    if (bytes <= w.cap) return 0;
    if (w.buf) {
        if (check_cuda(device_sync(), "workspace sync")) return 1;
        cudaFree(w.buf);
        w.buf = nullptr;
        w.cap = 0;
    }
    size_t cap = std::max(bytes + bytes / 4, (size_t)1 << 20);
    cap = (cap + 255) & ~(size_t)255;
    if (check_cuda(dev_malloc((void**)&w.buf, cap), "workspace cudaMalloc")) return 1;
    w.cap = cap;
    return 0;
}

DM view(const GridShape& g, const DM& a, int mc, int mr, i64 i0, i64 j0, i64 h, i64 w, int es) {
    const LocalShape L = local_shape(g, a, mc, mr);
    DM v = a;
    v.height = h;
    v.width = w;
    v.colAlign = (int)mod_(a.colAlign + i0, L.colStride);
    v.rowAlign = (int)mod_(a.rowAlign + j0, L.rowStride);
    const i64 iLoc = length_(i0, L.colShift, L.colStride);
    const i64 jLoc = length_(j0, L.rowShift, L.rowStride);
    v.buf = a.buf ? (char*)a.buf + (iLoc + jLoc * a.ldim) * es : nullptr;
    return v;
}

namespace {

struct Msg {
    int peer;
    Prog pa, pb;  // logical rows / cols of the message, in destination orientation
    i64 count() const { return pa.count * pb.count; }
    i64 off = 0;  // byte offset inside the send or recv region
};

inline void progs(const DM& src, bool transpose, const LocalShape& Ls, const LocalShape& Ld, Prog& pa, Prog& pb) {
    if (!transpose) {
        pa = intersect(src.height, Ls.colShift, Ls.colStride, Ld.colShift, Ld.colStride);
        pb = intersect(src.width, Ls.rowShift, Ls.rowStride, Ld.rowShift, Ld.rowStride);
    } else {
        pa = intersect(src.width, Ls.rowShift, Ls.rowStride, Ld.colShift, Ld.colStride);
        pb = intersect(src.height, Ls.colShift, Ls.colStride, Ld.rowShift, Ld.rowStride);
    }
}

template <typename T>
CopyDesc pack_desc(const DM& src, bool transpose, const Msg& m, void* out) {
    CopyDesc d;
    const T* base = static_cast<const T*>(src.buf);
    if (!transpose) {
        d.src = base + m.pa.src0 + m.pb.src0 * src.ldim;
        d.s_rs = m.pa.srcStep;
        d.s_cs = m.pb.srcStep * src.ldim;
    } else {  // logical rows run over the source's local columns
        d.src = base + m.pb.src0 + m.pa.src0 * src.ldim;
        d.s_rs = m.pa.srcStep * src.ldim;
        d.s_cs = m.pb.srcStep;
    }
    d.dst = out;
    d.d_rs = 1;
    d.d_cs = m.pa.count;
    d.nr = m.pa.count;
    d.nc = m.pb.count;
    return d;
}
template <typename T>
void dst_lattice(const DM& dst, const Msg& m, void*& p, i64& rs, i64& cs) {
    p = static_cast<T*>(dst.buf) + m.pa.dst0 + m.pb.dst0 * dst.ldim;
    rs = m.pa.dstStep;
    cs = m.pb.dstStep * dst.ldim;
}

// ctas_per_sm: grid cap.  Chunk-pipelined exchanges run the pack (+ push) of chunk q+1 and the unpack of chunk q at
// the same time; each gets half of the 4 resident CTAs per SM, so neither queues behind the other's grid.
template <typename T>
int launch_copies(const std::vector<CopyDesc>& v, cudaStream_t s, int ctas_per_sm = 8) {
    for (size_t i0 = 0; i0 < v.size(); i0 += kMaxDesc) {
        CopyBatch b;
        b.n = (int)std::min((size_t)kMaxDesc, v.size() - i0);
        i64 mx = 0;
        for (int i = 0; i < b.n; ++i) { b.d[i] = v[i0 + i]; mx = std::max(mx, b.d[i].nr * b.d[i].nc); }
        if (mx == 0) continue;
        const i64 want = (mx + 4095) / 4096;
        const i64 cap = std::max<i64>(1, (i64)num_sms() * ctas_per_sm / b.n);
        dim3 grid((unsigned)(std::max<i64>(1, std::min(want, cap)) * b.n));
        lattice_copy_kernel<T><<<grid, 256, 0, s>>>(b);
        ++g_launches;
        if (check_cuda(cudaGetLastError(), "lattice_copy_kernel")) return 1;
    }
    return 0;
}

template <typename T>
int launch_reduces(const std::vector<ReduceDesc>& v, T alpha, cudaStream_t s) {
    for (size_t i0 = 0; i0 < v.size(); i0 += kMaxRDesc) {
        ReduceBatch b;
        b.n = (int)std::min((size_t)kMaxRDesc, v.size() - i0);
        i64 mx = 0;
        for (int i = 0; i < b.n; ++i) { b.d[i] = v[i0 + i]; mx = std::max(mx, b.d[i].nr * b.d[i].nc); }
        if (mx == 0) continue;
        const i64 want = (mx + 1023) / 1024;
        const i64 cap = std::max<i64>(1, (i64)num_sms() * 8 / b.n);
        dim3 grid((unsigned)std::max<i64>(1, std::min(want, cap)), (unsigned)b.n);
        lattice_reduce_kernel<T><<<grid, 256, 0, s>>>(alpha, b);
        ++g_launches;
        if (check_cuda(cudaGetLastError(), "lattice_reduce_kernel")) return 1;
    }
    return 0;
}

inline i64 align16(i64 x) { return (x + 15) & ~(i64)15; }

int check_pair(const GridShape& sh, const DM& src, const DM& dst, bool transpose, const char* who) {
    if (!pair_supported(src.colDist, src.rowDist) || !pair_supported(dst.colDist, dst.rowDist))
        return set_error("%s: unsupported distribution pair (only MC/MR/VC/VR/STAR element distributions)", who);
    const i64 sh_ = transpose ? src.width : src.height, sw_ = transpose ? src.height : src.width;
    if (dst.height != sh_ || dst.width != sw_)
        return set_error("%s: nonconformal: src %lldx%lld%s dst %lldx%lld", who, src.height, src.width,
                         transpose ? "^T" : "", dst.height, dst.width);
    if (src.colAlign < 0 || src.colAlign >= sh.stride(src.colDist) || src.rowAlign < 0 ||
        src.rowAlign >= sh.stride(src.rowDist) || dst.colAlign < 0 || dst.colAlign >= sh.stride(dst.colDist) ||
        dst.rowAlign < 0 || dst.rowAlign >= sh.stride(dst.rowDist))
        return set_error("%s: alignment out of range", who);
    return 0;
}

}  // namespace

// Host-only: bytes rank `rank` sends to / receives from every owning-comm rank for dst := src (or src^T).
// Entry [rank] is the part that stays local.  Used by the planner tests (no GPU needed).
int redist_plan(const GridShape& sh, int rank, const DM& src, const DM& dst, bool transpose, int es, i64* sendBytes,
                i64* recvBytes) {
    if (check_pair(sh, src, dst, transpose, "hyd_redist_plan")) return 1;
    const int P = sh.size();
    int mc, mr;
    sh.coords(rank, mc, mr);
    bool mcFree, mrFree;
    free_coords(src.colDist, src.rowDist, mcFree, mrFree);
    const LocalShape Ls_me = local_shape(sh, src, mc, mr), Ld_me = local_shape(sh, dst, mc, mr);
    for (int p = 0; p < P; ++p) {
        sendBytes[p] = recvBytes[p] = 0;
        int pmc, pmr;
        sh.coords(p, pmc, pmr);
        if ((mcFree && pmc != mc) || (mrFree && pmr != mr)) continue;
        Prog pa, pb;
        progs(src, transpose, Ls_me, local_shape(sh, dst, pmc, pmr), pa, pb);
        sendBytes[p] = pa.count * pb.count * es;
        progs(src, transpose, local_shape(sh, src, pmc, pmr), Ld_me, pa, pb);
        recvBytes[p] = pa.count * pb.count * es;
    }
    return 0;
}

template <typename T>
int redistribute(Grid& g, const DM& src, const DM& dst, bool transpose, int slot, cudaStream_t s) {
    const GridShape& sh = g.shape;
    if (check_pair(sh, src, dst, transpose, "hyd_redist")) return 1;
    if (src.height == 0 || src.width == 0) return 0;
    NvtxRange nvtx(transpose ? "Transpose" : "Copy");
    ProfScope prof(g, Grid::PROF_REDIST, s);
    const int P = sh.size(), me = g.rank;
    bool mcFree, mrFree;
    free_coords(src.colDist, src.rowDist, mcFree, mrFree);
    std::vector<LocalShape> Ls(P), Ld(P);
    std::vector<int> pmc(P), pmr(P);
    for (int p = 0; p < P; ++p) {
        sh.coords(p, pmc[p], pmr[p]);
        Ls[p] = local_shape(sh, src, pmc[p], pmr[p]);
        Ld[p] = local_shape(sh, dst, pmc[p], pmr[p]);
    }
    if (Ls[me].lh > 0 && Ls[me].lw > 0 && (!src.buf || src.ldim < Ls[me].lh))
        return set_error("hyd_redist: bad source buffer / ldim");
    if (Ld[me].lh > 0 && Ld[me].lw > 0 && (!dst.buf || dst.ldim < Ld[me].lh))
        return set_error("hyd_redist: bad destination buffer / ldim");
    // data moves only between ranks that agree on the coordinates along which the source is replicated
    auto talks = [&](int q, int p) { return (!mcFree || pmc[q] == pmc[p]) && (!mrFree || pmr[q] == pmr[p]); };

    std::vector<Msg> sends, recvs;
    Msg self_msg;
    bool has_self = false;
    i64 send_bytes = 0, recv_bytes = 0;
    for (int p = 0; p < P; ++p) {
        if (!talks(me, p)) continue;
        Msg m;
        m.peer = p;
        if (p == me) {
            progs(src, transpose, Ls[me], Ld[me], m.pa, m.pb);
            if (m.count() > 0) { self_msg = m; has_self = true; }
            continue;
        }
        progs(src, transpose, Ls[me], Ld[p], m.pa, m.pb);
        if (m.count() > 0) { m.off = send_bytes; send_bytes = align16(send_bytes + m.count() * (i64)sizeof(T)); sends.push_back(m); }
        progs(src, transpose, Ls[p], Ld[me], m.pa, m.pb);
        if (m.count() > 0) { m.off = recv_bytes; recv_bytes = align16(recv_bytes + m.count() * (i64)sizeof(T)); recvs.push_back(m); }
    }

    const bool push = P > 1 && g.world->can_push();
    char* send_base = nullptr;   // staged path only
    char* recv_base = nullptr;
    std::vector<unsigned char> to(P, 0), from(P, 0);
    i64 need = 0, self_off = 0;
    if (push) {
        // Receive-region layout at EVERY rank (senders in ascending rank order), computed identically everywhere,
        // so that each sender knows where its message lands and all ranks agree on the region size.
        std::vector<i64> my_send_off(P, 0);
        for (int p = 0; p < P; ++p) {
            i64 run = 0;
            for (int q = 0; q < P; ++q) {
                if (q == p || !talks(q, p)) continue;
                Prog pa, pb;
                progs(src, transpose, Ls[q], Ld[p], pa, pb);
                const i64 bytes = pa.count * pb.count * (i64)sizeof(T);
                if (bytes == 0) continue;
                if (q == me) my_send_off[p] = run;
                if (p == me) for (Msg& m : recvs) if (m.peer == q) m.off = run;
                run = align16(run + bytes);
            }
            // room for p's own part behind its peers' messages (used by the merged unpack, see below)
            Prog pa, pb;
            progs(src, transpose, Ls[p], Ld[p], pa, pb);
            if (p == me) self_off = run;
            need = std::max(need, align16(run + pa.count * pb.count * (i64)sizeof(T)));
        }
        for (Msg& m : sends) { m.off = my_send_off[m.peer]; to[m.peer] = 1; }
        for (const Msg& m : recvs) from[m.peer] = 1;
        if (need > 0 && g.world->shared_reserve(slot, (size_t)need)) return 1;
        recv_base = g.world->shared_local(slot);
    } else {
        Workspace& w = g.ws[slot];
        if (send_bytes + recv_bytes > 0 && ws_reserve(w, (size_t)(send_bytes + recv_bytes))) return 1;
        send_base = w.buf;
        recv_base = w.buf + send_bytes;
    }

    // pack (+ the local part straight from source lattice to destination lattice).  On the push path the pack
    // kernel's destinations are the peers' receive regions: packing and the NVLink transfer are one kernel.
    // Large exchanges are split into K column chunks of every message: chunk q is unpacked on a side stream while
    // chunk q+1 is still being pushed, so the (HBM-bound) unpack and local part hide under the (NVLink-bound) push.
    int K = 1;
    if (push && need >= ((i64)48 << 20) && g.world->push_chunks()) K = (int)std::min<i64>(8, need / ((i64)24 << 20));
    auto chunk = [&](const Msg& m, int q, Msg& out) {  // columns [c0, c1) of message m
        const i64 c0 = m.pb.count * q / K, c1 = m.pb.count * (q + 1) / K;
        out = m;
        out.pb.count = c1 - c0;
        out.pb.src0 = m.pb.src0 + c0 * m.pb.srcStep;
        out.pb.dst0 = m.pb.dst0 + c0 * m.pb.dstStep;
        out.off = m.off + c0 * m.pa.count * (i64)sizeof(T);
        return out.pb.count > 0;
    };
    // Merged unpack (lattice_merge_kernel): the parts of my destination block -- the received messages and my own
    // part -- interleave its rows with period S = number of parts.  My own part is then packed like a message (into
    // my receive region) and ONE pass writes the block in full vectors instead of S strided partial-sector passes.
    std::vector<Msg> parts;
    bool merge = false;
    if (push && need > 0) {
        parts = recvs;
        if (has_self) { Msg m = self_msg; m.off = self_off; parts.push_back(m); }
        const i64 S = (i64)parts.size();
        if (S >= 2 && S <= kMaxMerge) {
            merge = true;
            i64 R = 0;
            std::vector<char> seen(S, 0);
            for (const Msg& m : parts) R += m.pa.count;
            for (const Msg& m : parts) {
                const Prog& b0 = parts[0].pb;
                if (m.pa.dstStep != S || m.pa.dst0 < 0 || m.pa.dst0 >= S || seen[m.pa.dst0] || m.pb.count != b0.count ||
                    m.pb.dst0 != b0.dst0 || m.pb.dstStep != b0.dstStep || m.pa.count != length_(R, m.pa.dst0, S)) {
                    merge = false;
                    break;
                }
                seen[m.pa.dst0] = 1;
            }
        }
    }
    // The rank's own part of a non-transposing merged redistribution is read by the merge kernel straight from the source
    // lattice (its rows are contiguous there): no pack, one read + one write less of that part (half of the block when two
    // grid rows feed a [*,VR] block).  A transposed own part keeps the tiled pack (its rows are ldim apart in the source).
    const bool self_direct = merge && has_self && !transpose && self_msg.pa.srcStep == 1;
    // Copy-engine push: a message whose rows are contiguous in the source (no transpose, the destination's row set is the
    // source's: [MC,MR] -> [*,VR] between the ranks of a grid column, a whole-block shift, ...) is a strided 2-D copy.  The copy
    // engines move it at the full peer-copy rate (SM stores reach ~80 % of it) and leave every SM to the merge / unpack
    // kernel of the previous chunk.  Only for large exchanges over real peer links (one process per GPU), all messages of
    // this rank qualifying, at most two peers (the copies of one rank are issued one after the other).  HYD_REDIST_CE=0 disables.
    static const bool ce_env = !(getenv("HYD_REDIST_CE") && atoi(getenv("HYD_REDIST_CE")) == 0);
    // (restricted to exchanges that end in the merged unpack -- the form measured and round-trip-verified on 8 GPUs; a plain
    // whole-block shift would qualify too but has not run on real peer links yet)
    bool ce_push = ce_env && push && merge && K > 1 && !transpose && !sends.empty() && sends.size() <= 2 && g.world->push_chunks();
    for (const Msg& m : sends) ce_push = ce_push && m.pa.srcStep == 1 && m.count() * (i64)sizeof(T) >= ((i64)8 << 20);
    static const int chunk_ctas = getenv("HYD_CHUNK_CTAS_PER_SM") ? std::max(1, atoi(getenv("HYD_CHUNK_CTAS_PER_SM"))) : 3;
    const int per_sm = K > 1 ? chunk_ctas : 8;
    cudaStream_t us = s;
    if (K > 1) {
        if (!g.unpack_stream) {
            int lo, hi;
            cudaDeviceGetStreamPriorityRange(&lo, &hi);
            if (check_cuda(cudaStreamCreateWithPriority(&g.unpack_stream, cudaStreamNonBlocking, hi), "unpack stream")) return 1;
            cudaEventCreateWithFlags(&g.ev_unpack_fork, cudaEventDisableTiming);
            cudaEventCreateWithFlags(&g.ev_unpack_join, cudaEventDisableTiming);
            for (cudaEvent_t& e : g.ev_chunk) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        }
        us = g.unpack_stream;
        if (check_cuda(cudaEventRecord(g.ev_unpack_fork, s), "unpack fork")) return 1;
        cudaStreamWaitEvent(us, g.ev_unpack_fork, 0);
    }
    std::vector<CopyDesc> descs;
    if (push && need > 0 && g.world->push_open(slot, to.data(), s)) return 1;
    for (int q = 0; q < K; ++q) {
        if (q > 0) g.world->push_next(slot);
        descs.clear();
        Msg c;
        for (const Msg& m : sends) {
            if (!chunk(m, q, c)) continue;
            if (ce_push) {
                // whole (strided) columns of the source: one 2-D copy-engine transfer straight into the peer's region
                const char* sp = static_cast<const char*>(src.buf) + (c.pa.src0 + c.pb.src0 * src.ldim) * (i64)sizeof(T);
                ProfScope pk(g, Grid::PROF_PACK, s);
                if (check_cuda(cudaMemcpy2DAsync(g.world->shared_peer(slot, c.peer) + c.off, (size_t)c.pa.count * sizeof(T), sp,
                                                 (size_t)(c.pb.srcStep * src.ldim) * sizeof(T), (size_t)c.pa.count * sizeof(T),
                                                 (size_t)c.pb.count, cudaMemcpyDeviceToDevice, s),
                               "hyd_redist: copy-engine push"))
                    return 1;
                continue;
            }
            descs.push_back(pack_desc<T>(src, transpose, c, push ? g.world->shared_peer(slot, c.peer) + c.off : send_base + c.off));
        }
        if (has_self && merge && self_direct) {
            // nothing to pack: see the merge descriptor below
        } else if (has_self && merge) {
            Msg m = self_msg;
            m.off = self_off;
            if (chunk(m, q, c)) descs.push_back(pack_desc<T>(src, transpose, c, recv_base + c.off));
        } else if (has_self && chunk(self_msg, q, c)) {
            CopyDesc d = pack_desc<T>(src, transpose, c, nullptr);
            dst_lattice<T>(dst, c, d.dst, d.d_rs, d.d_cs);
            descs.push_back(d);
        }
        {
            ProfScope pk(g, Grid::PROF_PACK, s);
            if (launch_copies<T>(descs, s, per_sm)) return 1;
        }

        if (push) {
            if (need > 0 && (g.world->push_post(slot, to.data(), s) || g.world->push_wait(slot, from.data(), us))) return 1;
        } else if (!sends.empty() || !recvs.empty() || P > 1) {
            std::vector<i64> sc(P, 0), so(P, 0), rc(P, 0), ro(P, 0);
            for (const Msg& m : sends) { sc[m.peer] = m.count() * (i64)sizeof(T); so[m.peer] = m.off; }
            for (const Msg& m : recvs) { rc[m.peer] = m.count() * (i64)sizeof(T); ro[m.peer] = m.off; }
            if (g.world->exchange(send_base, sc.data(), so.data(), recv_base, rc.data(), ro.data(), s)) return 1;
        }
        if (merge) {
            if (us != s) {  // my own part of this chunk was packed on s
                if (check_cuda(cudaEventRecord(g.ev_chunk[q], s), "chunk event")) return 1;
                cudaStreamWaitEvent(us, g.ev_chunk[q], 0);
            }
            MergeDesc md;
            md.S = (int)parts.size();
            md.nr = 0;
            md.nc = 0;
            for (size_t pi = 0; pi < parts.size(); ++pi) {
                const Msg& m = parts[pi];
                md.nr += m.pa.count;
                if (!chunk(m, q, c)) continue;
                if (self_direct && pi + 1 == parts.size()) {   // my own part (always the last one): a lattice of the source
                    md.src[m.pa.dst0] = static_cast<const T*>(src.buf) + c.pa.src0 + c.pb.src0 * src.ldim;
                    md.s_rs[m.pa.dst0] = c.pa.srcStep;
                    md.s_cs[m.pa.dst0] = c.pb.srcStep * src.ldim;
                } else {
                    md.src[m.pa.dst0] = recv_base + c.off;
                    md.s_rs[m.pa.dst0] = 1;
                    md.s_cs[m.pa.dst0] = c.pa.count;
                }
                md.nc = c.pb.count;
                md.dst = static_cast<T*>(dst.buf) + c.pb.dst0 * dst.ldim;
                md.d_cs = c.pb.dstStep * dst.ldim;
            }
            if (md.nc > 0 && md.nr > 0) {
                ProfScope pu(g, Grid::PROF_UNPACK, us);
                const i64 V = 16 / (i64)sizeof(T);
                const i64 want = ((md.nr + V - 1) / V + 255) / 256 * ((md.nc + 7) / 8);
                const unsigned grid = (unsigned)std::max<i64>(1, std::min<i64>(want, (i64)num_sms() * per_sm));
                lattice_merge_kernel<T><<<grid, 256, 0, us>>>(md);
                ++g_launches;
                if (check_cuda(cudaGetLastError(), "lattice_merge_kernel")) return 1;
            }
            continue;
        }
        descs.clear();
        for (const Msg& m : recvs) {
            if (!chunk(m, q, c)) continue;
            CopyDesc d;
            d.src = recv_base + c.off;
            d.s_rs = 1;
            d.s_cs = c.pa.count;
            dst_lattice<T>(dst, c, d.dst, d.d_rs, d.d_cs);
            d.nr = c.pa.count;
            d.nc = c.pb.count;
            descs.push_back(d);
        }
        {
            ProfScope pu(g, Grid::PROF_UNPACK, us);
            if (launch_copies<T>(descs, us, per_sm)) return 1;
        }
    }
    if (K > 1) {
        if (check_cuda(cudaEventRecord(g.ev_unpack_join, us), "unpack join")) return 1;
        cudaStreamWaitEvent(s, g.ev_unpack_join, 0);
    }
    if (push && need > 0) return g.world->push_close(slot, s);
    return 0;
}

template <typename T>
int redistribute_direct(Grid& g, const DM& src, DM& dst, bool transpose, int slot, i64 alloc_width, cudaStream_t s) {
    const GridShape& sh = g.shape;
    if (check_pair(sh, src, dst, transpose, "hyd_redist(direct)")) return 1;
    ProfScope prof(g, Grid::PROF_REDIST, s);
    const int P = sh.size(), me = g.rank;
    bool mcFree, mrFree;
    free_coords(src.colDist, src.rowDist, mcFree, mrFree);
    std::vector<LocalShape> Ls(P), Ld(P);
    std::vector<int> pmc(P), pmr(P);
    std::vector<i64> ldp(P);
    const i64 q16 = 16 / (i64)sizeof(T);
    i64 need = 0;
    for (int p = 0; p < P; ++p) {
        sh.coords(p, pmc[p], pmr[p]);
        Ls[p] = local_shape(sh, src, pmc[p], pmr[p]);
        Ld[p] = local_shape(sh, dst, pmc[p], pmr[p]);
        ldp[p] = std::max<i64>(q16, (Ld[p].lh + q16 - 1) / q16 * q16);
        need = std::max(need, ldp[p] * std::max<i64>(alloc_width, 1) * (i64)sizeof(T));
    }
    if (g.world->shared_reserve(slot, (size_t)need)) return 1;
    dst.buf = g.world->shared_local(slot);
    dst.ldim = ldp[me];
    if (Ls[me].lh > 0 && Ls[me].lw > 0 && (!src.buf || src.ldim < Ls[me].lh))
        return set_error("hyd_redist(direct): bad source buffer / ldim");
    auto talks = [&](int q, int p) { return (!mcFree || pmc[q] == pmc[p]) && (!mrFree || pmr[q] == pmr[p]); };
    std::vector<unsigned char> to(P, 0), from(P, 0);
    std::vector<CopyDesc> descs;
    for (int p = 0; p < P; ++p) {
        if (!talks(me, p)) continue;
        Msg m;
        m.peer = p;
        progs(src, transpose, Ls[me], Ld[p], m.pa, m.pb);
        if (m.count() > 0) {
            CopyDesc d = pack_desc<T>(src, transpose, m, nullptr);
            d.dst = reinterpret_cast<T*>(g.world->shared_peer(slot, p)) + m.pa.dst0 + m.pb.dst0 * ldp[p];
            d.d_rs = m.pa.dstStep;
            d.d_cs = m.pb.dstStep * ldp[p];
            descs.push_back(d);
            if (p != me) to[p] = 1;
        }
        if (p != me) {
            Prog pa, pb;
            progs(src, transpose, Ls[p], Ld[me], pa, pb);
            if (pa.count * pb.count > 0) from[p] = 1;
        }
    }
    if (g.world->push_open(slot, to.data(), s)) return 1;
    if (launch_copies<T>(descs, s)) return 1;
    if (g.world->push_post(slot, to.data(), s)) return 1;
    return g.world->push_wait(slot, from.data(), s);
}

template <typename T>
int reduce_into(Grid& g, T alpha, const DM& src, const DM& dst, bool transpose, int slot, cudaStream_t s) {
    const GridShape& sh = g.shape;
    if (check_pair(sh, src, dst, transpose, "hyd_axpy_contract")) return 1;
    if (src.height == 0 || src.width == 0) return 0;
    NvtxRange nvtx("AxpyContract");
    ProfScope prof(g, Grid::PROF_REDUCE, s);
    const int P = sh.size();
    bool mcFree, mrFree;
    free_coords(src.colDist, src.rowDist, mcFree, mrFree);
    const LocalShape Ls_me = local_shape(sh, src, g.mc, g.mr), Ld_me = local_shape(sh, dst, g.mc, g.mr);
    if (Ls_me.lh > 0 && Ls_me.lw > 0 && (!src.buf || src.ldim < Ls_me.lh))
        return set_error("hyd_axpy_contract: bad source buffer / ldim");
    if (Ld_me.lh > 0 && Ld_me.lw > 0 && (!dst.buf || dst.ldim < Ld_me.lh))
        return set_error("hyd_axpy_contract: bad destination buffer / ldim");

    struct Group { int kmc, kmr; Msg m; std::vector<i64> member_off; };
    std::vector<Msg> sends, recvs;
    std::vector<Group> groups;
    i64 send_bytes = 0, recv_bytes = 0, self_off = -1;
    Msg self_msg;
    // senders: every rank contributes its partial sum to every destination that owns the element
    for (int p = 0; p < P; ++p) {
        if (p == g.rank) continue;
        int pmc, pmr;
        sh.coords(p, pmc, pmr);
        Msg m;
        m.peer = p;
        progs(src, transpose, Ls_me, local_shape(sh, dst, pmc, pmr), m.pa, m.pb);
        if (m.count() > 0) { m.off = send_bytes; send_bytes = align16(send_bytes + m.count() * (i64)sizeof(T)); sends.push_back(m); }
    }
    // receivers: ascending VC rank fixes the summation order (the oracle uses the same order)
    for (int v = 0; v < P; ++v) {
        const int pmc = v % sh.r, pmr = v / sh.r;
        const int p = sh.rank_of(pmc, pmr);
        Msg m;
        m.peer = p;
        progs(src, transpose, local_shape(sh, src, pmc, pmr), Ld_me, m.pa, m.pb);
        if (m.count() == 0) continue;
        m.off = recv_bytes;
        recv_bytes = align16(recv_bytes + m.count() * (i64)sizeof(T));
        recvs.push_back(m);
        if (p == g.rank) { self_off = m.off; self_msg = m; }
        const int kmc = mcFree ? -1 : pmc, kmr = mrFree ? -1 : pmr;
        Group* grp = nullptr;
        for (auto& q : groups) if (q.kmc == kmc && q.kmr == kmr) grp = &q;
        if (!grp) { groups.push_back(Group{kmc, kmr, m, {}}); grp = &groups.back(); }
        grp->member_off.push_back(m.off);
    }
    const bool push = P > 1 && g.world->can_push();
    char* send_base = nullptr;
    char* recv_base = nullptr;
    std::vector<unsigned char> to(P, 0), from(P, 0);
    i64 need = 0;
    if (push) {
        // receive-region layout at every rank: members in ascending VC rank (the summation order), self included --
        // exactly the order the recvs / member offsets above were built in, so only the senders' view is new here
        std::vector<i64> my_send_off(P, 0);
        for (int p = 0; p < P; ++p) {
            int dmc, dmr;
            sh.coords(p, dmc, dmr);
            const LocalShape Ldp = local_shape(sh, dst, dmc, dmr);
            i64 run = 0;
            for (int v = 0; v < P; ++v) {
                const int qmc = v % sh.r, qmr = v / sh.r;
                const int q = sh.rank_of(qmc, qmr);
                Prog pa, pb;
                progs(src, transpose, local_shape(sh, src, qmc, qmr), Ldp, pa, pb);
                const i64 bytes = pa.count * pb.count * (i64)sizeof(T);
                if (bytes == 0) continue;
                if (q == g.rank) my_send_off[p] = run;
                run = align16(run + bytes);
            }
            need = std::max(need, run);
        }
        for (Msg& m : sends) { m.off = my_send_off[m.peer]; to[m.peer] = 1; }
        for (const Msg& m : recvs) if (m.peer != g.rank) from[m.peer] = 1;
        if (need > 0 && g.world->shared_reserve(slot, (size_t)need)) return 1;
        recv_base = g.world->shared_local(slot);
    } else {
        Workspace& w = g.ws[slot];
        if (send_bytes + recv_bytes > 0 && ws_reserve(w, (size_t)(send_bytes + recv_bytes))) return 1;
        send_base = w.buf;
        recv_base = w.buf + send_bytes;
    }

    std::vector<CopyDesc> descs;
    if (push && need > 0 && g.world->push_open(slot, to.data(), s)) return 1;
    for (const Msg& m : sends)
        descs.push_back(pack_desc<T>(src, transpose, m, push ? g.world->shared_peer(slot, m.peer) + m.off : send_base + m.off));
    if (self_off >= 0) descs.push_back(pack_desc<T>(src, transpose, self_msg, recv_base + self_off));
    if (launch_copies<T>(descs, s)) return 1;

    if (push) {
        if (need > 0 && (g.world->push_post(slot, to.data(), s) || g.world->push_wait(slot, from.data(), s))) return 1;
    } else if (P > 1) {
        std::vector<i64> sc(P, 0), so(P, 0), rc(P, 0), ro(P, 0);
        for (const Msg& m : sends) { sc[m.peer] = m.count() * (i64)sizeof(T); so[m.peer] = m.off; }
        for (const Msg& m : recvs) if (m.peer != g.rank) { rc[m.peer] = m.count() * (i64)sizeof(T); ro[m.peer] = m.off; }
        if (g.world->exchange(send_base, sc.data(), so.data(), recv_base, rc.data(), ro.data(), s)) return 1;
    }

    std::vector<ReduceDesc> rdescs;
    for (const Group& q : groups) {
        for (size_t m0 = 0; m0 < q.member_off.size(); m0 += kMaxMembers) {
            ReduceDesc d;
            i64 rs, cs;
            dst_lattice<T>(dst, q.m, d.dst, rs, cs);
            d.d_rs = rs;
            d.d_cs = cs;
            d.nr = q.m.pa.count;
            d.nc = q.m.pb.count;
            d.nq = (int)std::min((size_t)kMaxMembers, q.member_off.size() - m0);
            for (int i = 0; i < d.nq; ++i) d.member[i] = recv_base + q.member_off[m0 + i];
            rdescs.push_back(d);
        }
    }
    // groups touch disjoint destination lattices, but chunks of one group do not: launch chunk by chunk
    bool chunked = false;
    for (const Group& q : groups) if (q.member_off.size() > (size_t)kMaxMembers) chunked = true;
    if (!chunked) {
        if (launch_reduces<T>(rdescs, alpha, s)) return 1;
    } else {
        for (const ReduceDesc& d : rdescs) {
            std::vector<ReduceDesc> one(1, d);
            if (launch_reduces<T>(one, alpha, s)) return 1;
        }
    }
    if (push && need > 0) return g.world->push_close(slot, s);
    return 0;
}

template int redistribute<double>(Grid&, const DM&, const DM&, bool, int, cudaStream_t);
template int redistribute<float>(Grid&, const DM&, const DM&, bool, int, cudaStream_t);
template int redistribute<h16>(Grid&, const DM&, const DM&, bool, int, cudaStream_t);
template int redistribute_direct<double>(Grid&, const DM&, DM&, bool, int, i64, cudaStream_t);
template int redistribute_direct<float>(Grid&, const DM&, DM&, bool, int, i64, cudaStream_t);
template int reduce_into<double>(Grid&, double, const DM&, const DM&, bool, int, cudaStream_t);
template int reduce_into<float>(Grid&, float, const DM&, const DM&, bool, int, cudaStream_t);

void grid_destroy(Grid* g) {
    if (!g) return;
    cudaDeviceSynchronize();
    for (auto& w : g->ws) if (w.buf) cudaFree(w.buf);
    for (auto& w : g->panel) if (w.buf) cudaFree(w.buf);
    for (auto& w : g->host_stage) if (w.buf) cudaFree(w.buf);
    if (g->coll_ws.buf) cudaFree(g->coll_ws.buf);
    if (g->level1_ws.buf) cudaFree(g->level1_ws.buf);
    if (g->trrk_ws.buf) cudaFree(g->trrk_ws.buf);
    for (auto& b : g->pool) {
        if (b.w.buf) cudaFree(b.w.buf);
        if (b.released) cudaEventDestroy(b.released);
    }
    for (auto& v : g->prof_ev) for (cudaEvent_t e : v) cudaEventDestroy(e);
    for (cudaEvent_t e : g->upload_ev) cudaEventDestroy(e);
    for (int i = 0; i < 2; ++i) {
        if (g->comm_stream[i]) cudaStreamDestroy(g->comm_stream[i]);
        if (g->ev_panel_ready[i]) cudaEventDestroy(g->ev_panel_ready[i]);
        if (g->ev_panel_free[i]) cudaEventDestroy(g->ev_panel_free[i]);
    }
    if (g->ev_fork) cudaEventDestroy(g->ev_fork);
    for (cudaEvent_t e : g->ev_pipe) if (e) cudaEventDestroy(e);
    if (g->ev_join) cudaEventDestroy(g->ev_join);
    if (g->unpack_stream) cudaStreamDestroy(g->unpack_stream);
    if (g->download_stream) cudaStreamDestroy(g->download_stream);
    if (g->ev_block_done) cudaEventDestroy(g->ev_block_done);
    if (g->ev_download_done) cudaEventDestroy(g->ev_download_done);
    if (g->ev_unpack_fork) cudaEventDestroy(g->ev_unpack_fork);
    if (g->ev_unpack_join) cudaEventDestroy(g->ev_unpack_join);
    for (cudaEvent_t e : g->ev_chunk) if (e) cudaEventDestroy(e);
    if (g->owns_world) delete g->world;
    delete g;
}

}  // namespace hyd
