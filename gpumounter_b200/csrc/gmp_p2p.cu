// Peer-to-peer probe: capability matrices (CUDA vs NVML) and NVLink 5 bandwidth with the same
// pattern kernels pointed at PEER memory.  Section 5 of include/gpumounter_probe.h.
//
// The exchange step needs no collective library: with peer access enabled, a kernel running on
// GPU i dereferences GPU j's buffer directly, so the 16/32-byte vector stores (or loads) of
// gmp_kernels.cu travel over NVLink/NVSwitch and the verify pass is fused with the transfer
// (writer on i stores, owner j verifies; reader on i loads and verifies in one pass).
// Bound: NVLink 5, 900 GB/s per direction per GPU nominal (18 links x 50 GB/s), uniform to every
// peer through NVSwitch - every pair should measure the same, an outlier is a finding.

#include "gmp_internal.h"

#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <thread>
#include <vector>

namespace {

using namespace gmp;

// 256 CTAs x 256 threads, every thread adds (its global index + 1) to a 64-bit counter that lives on the PEER:
// the adds are resolved by the remote GPU's L2 atomic unit after crossing NVLink.  Expected sum = n(n+1)/2.
constexpr unsigned kAtomicCtas = 256, kAtomicThreads = 256;
__global__ void __launch_bounds__(256) k_peer_atomic_add(unsigned long long* counter) {
    const unsigned long long mine = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x + 1;
    atomicAdd(counter, mine);
}

struct PeerGpu {
    int ordinal = -1;
    cudaStream_t stream = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    void* buf = nullptr;
    unsigned long long* acc = nullptr;
    bool touched = false;
    std::vector<int> enabled_peers;  // ordinals whose peer access THIS probe enabled (not those that already were)
};

struct Session {
    std::vector<PeerGpu> g;
    int prev_device = -1;
    bool keep = false;

    void teardown() {
        for (PeerGpu& p : g) {
            if (!p.touched) continue;
            if (cudaSetDevice(p.ordinal) != cudaSuccess) continue;
            if (p.acc) cudaFree(p.acc);
            if (p.buf) cudaFree(p.buf);
            if (p.e0) cudaEventDestroy(p.e0);
            if (p.e1) cudaEventDestroy(p.e1);
            if (p.stream) cudaStreamDestroy(p.stream);
            // no peer mapping is left behind, also when the contexts are kept: every later cudaMalloc of the process
            // would otherwise be mapped into all peers (measured: 2 x 1 GiB per GPU went from 2 ms to 60 ms per GPU)
            for (int peer : p.enabled_peers) (void)cudaDeviceDisablePeerAccess(peer);
            const int ordinal = p.ordinal;
            p = PeerGpu{};
            p.ordinal = ordinal;
            if (!keep) {
                (void)cudaDeviceReset();
                forget_device_state(p.ordinal);
            }
        }
        (void)cudaGetLastError();
        if (prev_device >= 0) (void)cudaSetDevice(prev_device);
    }
};

inline int cell(int n, int i, int j) { return i * n + j; }

uint64_t pair_seed(uint64_t base, int i, int j, uint32_t it) { return base + ((uint64_t)(i * 64 + j) << 20) + it; }

int fetch(PeerGpu& p, gmp_verify_result* v) {
    unsigned long long raw[4];
    GMP_CUDA(cudaMemcpyAsync(raw, p.acc, sizeof(raw), cudaMemcpyDeviceToHost, p.stream));
    GMP_CUDA(cudaStreamSynchronize(p.stream));
    decode_acc(raw, v);
    return GMP_OK;
}

void note(gmp_p2p_result* out, const gmp_verify_result& v, int src, int dst) {
    if (v.mismatch_lanes == 0) return;
    if (out->mismatch_lanes == 0) {
        out->first_bad_src = src;
        out->first_bad_dst = dst;
    }
    out->mismatch_lanes += v.mismatch_lanes;
}

// Owner j checks its own buffer against what i was supposed to have written.
int verify_local(Session& s, gmp_p2p_result* out, int i, int j, size_t off, size_t bytes, uint64_t seed) {
    PeerGpu& pj = s.g[(size_t)j];
    GMP_CUDA(cudaSetDevice(pj.ordinal));
    GMP_CUDA(launch_verify(static_cast<char*>(pj.buf) + off, bytes, GMP_PAT_SPLITMIX, seed, pj.acc, pj.stream));
    gmp_verify_result v;
    int rc = fetch(pj, &v);
    if (rc != GMP_OK) return rc;
    note(out, v, i, j);
    return GMP_OK;
}

int run_pairs(Session& s, const gmp_p2p_opts& o, gmp_p2p_result* out) {
    const int n = out->n;
    const LaunchCfg kPeerStoreCfg = peer_store_cfg(), kPeerLoadCfg = peer_load_cfg();
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) {
            if (i == j || !out->can_access[cell(n, i, j)]) continue;
            PeerGpu& pi = s.g[(size_t)i];
            PeerGpu& pj = s.g[(size_t)j];
            double best_w = 0.0, best_r = 0.0;
            for (uint32_t it = 0; it < o.iters; ++it) {
                const uint64_t seed = pair_seed(o.seed, i, j, it);
                float ms = 0.f;
                // i stores into j's HBM
                GMP_CUDA(cudaSetDevice(pi.ordinal));
                GMP_CUDA(cudaEventRecord(pi.e0, pi.stream));
                GMP_CUDA(launch_write(pj.buf, o.bytes, GMP_PAT_SPLITMIX, seed, pi.stream, &kPeerStoreCfg));
                GMP_CUDA(cudaEventRecord(pi.e1, pi.stream));
                GMP_CUDA(cudaEventSynchronize(pi.e1));
                GMP_CUDA(cudaEventElapsedTime(&ms, pi.e0, pi.e1));
                if (ms > 0.f) best_w = std::max(best_w, (double)o.bytes / (ms * 1e-3) / 1e9);
                int rc = verify_local(s, out, i, j, 0, o.bytes, seed);
                if (rc != GMP_OK) return rc;
                // i loads j's HBM and verifies in the same pass
                GMP_CUDA(cudaSetDevice(pi.ordinal));
                GMP_CUDA(cudaEventRecord(pi.e0, pi.stream));
                GMP_CUDA(launch_verify(pj.buf, o.bytes, GMP_PAT_SPLITMIX, seed, pi.acc, pi.stream, &kPeerLoadCfg));
                GMP_CUDA(cudaEventRecord(pi.e1, pi.stream));
                GMP_CUDA(cudaEventSynchronize(pi.e1));
                GMP_CUDA(cudaEventElapsedTime(&ms, pi.e0, pi.e1));
                if (ms > 0.f) best_r = std::max(best_r, (double)o.bytes / (ms * 1e-3) / 1e9);
                gmp_verify_result v;
                rc = fetch(pi, &v);
                if (rc != GMP_OK) return rc;
                note(out, v, j, i);
            }
            out->write_gbs[cell(n, i, j)] = best_w;
            out->read_gbs[cell(n, i, j)] = best_r;
            // library baseline: the copy engines move i's buffer into j's; j verifies what arrived
            double best_c = 0.0;
            for (uint32_t it = 0; it < o.iters; ++it) {
                const uint64_t seed = pair_seed(o.seed ^ 0xC0FFEEull, i, j, it);
                float ms = 0.f;
                GMP_CUDA(cudaSetDevice(pi.ordinal));
                GMP_CUDA(launch_write(pi.buf, o.bytes, GMP_PAT_SPLITMIX, seed, pi.stream));
                GMP_CUDA(cudaEventRecord(pi.e0, pi.stream));
                GMP_CUDA(cudaMemcpyPeerAsync(pj.buf, pj.ordinal, pi.buf, pi.ordinal, o.bytes, pi.stream));
                GMP_CUDA(cudaEventRecord(pi.e1, pi.stream));
                GMP_CUDA(cudaEventSynchronize(pi.e1));
                GMP_CUDA(cudaEventElapsedTime(&ms, pi.e0, pi.e1));
                if (ms > 0.f) best_c = std::max(best_c, (double)o.bytes / (ms * 1e-3) / 1e9);
                int rc = verify_local(s, out, i, j, 0, o.bytes, seed);
                if (rc != GMP_OK) return rc;
            }
            out->memcpy_gbs[cell(n, i, j)] = best_c;
            // native atomics from i on a counter in j's memory
            {
                unsigned long long got = 0;
                const unsigned long long total = (unsigned long long)kAtomicCtas * kAtomicThreads;
                GMP_CUDA(cudaSetDevice(pj.ordinal));
                GMP_CUDA(cudaMemsetAsync(pj.acc, 0, sizeof(unsigned long long), pj.stream));
                GMP_CUDA(cudaStreamSynchronize(pj.stream));
                GMP_CUDA(cudaSetDevice(pi.ordinal));
                k_peer_atomic_add<<<kAtomicCtas, kAtomicThreads, 0, pi.stream>>>(pj.acc);
                GMP_CUDA(cudaGetLastError());
                GMP_CUDA(cudaStreamSynchronize(pi.stream));
                GMP_CUDA(cudaSetDevice(pj.ordinal));
                GMP_CUDA(cudaMemcpyAsync(&got, pj.acc, sizeof(got), cudaMemcpyDeviceToHost, pj.stream));
                GMP_CUDA(cudaStreamSynchronize(pj.stream));
                const bool ok = got == total * (total + 1) / 2;
                out->atomics_ok[cell(n, i, j)] = ok ? 1 : 0;
                if (!ok) note(out, gmp_verify_result{1, 0, 0, 0}, i, j);
            }
        }
    }
    return GMP_OK;
}

// Round r of the circle-method 1-factorisation of K_m (m even): partner of x.
int partner_in_round(int m, int r, int x) {
    if (x == m - 1) return r;                 // the fixed element plays r
    if (x == r) return m - 1;
    return ((2 * r - x) % (m - 1) + (m - 1)) % (m - 1);
}

int run_bidir(Session& s, const gmp_p2p_opts& o, gmp_p2p_result* out) {
    const int n = out->n;
    const LaunchCfg kPeerStoreCfg = peer_store_cfg();
    const int m = (n % 2 == 0) ? n : n + 1;  // odd n: one GPU sits out each round
    std::vector<double> worst((size_t)n, 0.0);
    std::vector<bool> seen((size_t)n, false);
    for (int r = 0; r < m - 1; ++r) {
        std::vector<int> partner((size_t)n, -1);
        for (int x = 0; x < n; ++x) {
            const int p = partner_in_round(m, r, x);
            if (p < n && p != x && out->can_access[cell(n, x, p)] && out->can_access[cell(n, p, x)]) partner[(size_t)x] = p;
        }
        const uint64_t seed_base = o.seed + 0x1000000ull * (uint64_t)(r + 1);
        // queue every GPU's store kernel first, then wait: both directions of every pair are in flight together
        for (int x = 0; x < n; ++x) {
            const int p = partner[(size_t)x];
            if (p < 0) continue;
            PeerGpu& px = s.g[(size_t)x];
            GMP_CUDA(cudaSetDevice(px.ordinal));
            GMP_CUDA(cudaEventRecord(px.e0, px.stream));
            GMP_CUDA(launch_write(s.g[(size_t)p].buf, o.bytes, GMP_PAT_SPLITMIX, pair_seed(seed_base, x, p, 0), px.stream,
                                  &kPeerStoreCfg));
            GMP_CUDA(cudaEventRecord(px.e1, px.stream));
        }
        std::vector<float> ms((size_t)n, 0.f);
        for (int x = 0; x < n; ++x) {
            if (partner[(size_t)x] < 0) continue;
            PeerGpu& px = s.g[(size_t)x];
            GMP_CUDA(cudaSetDevice(px.ordinal));
            GMP_CUDA(cudaEventSynchronize(px.e1));
            GMP_CUDA(cudaEventElapsedTime(&ms[(size_t)x], px.e0, px.e1));
        }
        for (int x = 0; x < n; ++x) {
            const int p = partner[(size_t)x];
            if (p < 0) continue;
            const float t = std::max(ms[(size_t)x], ms[(size_t)p]);
            if (t > 0.f) {
                const double gbs = 2.0 * (double)o.bytes / (t * 1e-3) / 1e9;  // sent + received by x
                worst[(size_t)x] = seen[(size_t)x] ? std::min(worst[(size_t)x], gbs) : gbs;
                seen[(size_t)x] = true;
            }
            int rc = verify_local(s, out, p, x, 0, o.bytes, pair_seed(seed_base, p, x, 0));
            if (rc != GMP_OK) return rc;
        }
    }
    for (int x = 0; x < n; ++x) out->bidir_gbs[x] = worst[(size_t)x];
    return GMP_OK;
}

int run_alltoall(Session& s, const gmp_p2p_opts& o, gmp_p2p_result* out) {
    const int n = out->n;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j)
            if (i != j && !out->can_access[cell(n, i, j)]) return GMP_OK;  // not a full mesh: skip
    // buffer j is cut into n slices; GPU i owns slice i of every peer's buffer
    const size_t slice = ((size_t)o.bytes / (size_t)n) & ~(size_t)31;
    if (slice == 0) return GMP_OK;
    const uint64_t seed_base = o.seed + 0x7000000ull;
    float best_round = 0.f;  // best of o.iters rounds; a round lasts as long as its slowest sender
    for (uint32_t it = 0; it < o.iters; ++it) {
        for (int i = 0; i < n; ++i) {
            // ONE launch per sender: tile w goes to peer (i + 1 + w mod (n-1)) mod n, so every route of the sender carries
            // traffic for the whole kernel and no two senders aim at the same receiver at the same moment
            PeerGpu& pi = s.g[(size_t)i];
            void* dsts[GMP_MAX_DEVICES];
            uint64_t seeds[GMP_MAX_DEVICES];
            for (int k = 1; k < n; ++k) {
                const int j = (i + k) % n;
                dsts[k - 1] = static_cast<char*>(s.g[(size_t)j].buf) + (size_t)i * slice;
                seeds[k - 1] = pair_seed(seed_base, i, j, 0);
            }
            GMP_CUDA(cudaSetDevice(pi.ordinal));
            GMP_CUDA(cudaEventRecord(pi.e0, pi.stream));
            GMP_CUDA(launch_alltoall_write(dsts, seeds, n - 1, slice, pi.stream));
            GMP_CUDA(cudaEventRecord(pi.e1, pi.stream));
        }
        float slowest = 0.f;
        for (int i = 0; i < n; ++i) {
            PeerGpu& pi = s.g[(size_t)i];
            float ms = 0.f;
            GMP_CUDA(cudaSetDevice(pi.ordinal));
            GMP_CUDA(cudaEventSynchronize(pi.e1));
            GMP_CUDA(cudaEventElapsedTime(&ms, pi.e0, pi.e1));
            slowest = std::max(slowest, ms);
        }
        if (slowest > 0.f && (best_round == 0.f || slowest < best_round)) best_round = slowest;
    }
    const float slowest = best_round;
    if (slowest > 0.f) out->alltoall_gbs = (double)slice * n * (n - 1) / (slowest * 1e-3) / 1e9;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            if (i == j) continue;
            int rc = verify_local(s, out, i, j, (size_t)i * slice, slice, pair_seed(seed_base, i, j, 0));
            if (rc != GMP_OK) return rc;
        }
    return GMP_OK;
}

int fill_nvml_matrices(const char* const* uuids, gmp_p2p_result* out) {
    const int n = out->n;
    std::vector<int32_t> tmp[4];
    int32_t* ptr[4];
    for (int c = 0; c < 4; ++c) {
        tmp[c].assign((size_t)n * (size_t)n, 0);
        ptr[c] = tmp[c].data();
    }
    int32_t links[GMP_MAX_DEVICES];
    int rc = nvml_p2p_everything(uuids, n, ptr, links);
    if (rc != GMP_OK) return rc;
    int32_t* dst[4] = {out->nvml_read, out->nvml_write, out->nvml_nvlink, out->nvml_atomics};
    for (int c = 0; c < 4; ++c)
        for (int k = 0; k < n * n; ++k) dst[c][k] = tmp[c][(size_t)k];
    for (int i = 0; i < n; ++i) out->nvlink_active[i] = links[i];
    return GMP_OK;
}

int run(const char* const* uuids, const gmp_p2p_opts& o, gmp_p2p_result* out, Session& s) {
    const int n = out->n;
    uint64_t t0 = now_ns();
    int count = 0;
    for (int i = 0; i < n; ++i) {
        int ordinal = -1;
        int rc = find_cuda_device(uuids[i], &ordinal, &count);
        if (rc != GMP_OK) return rc;
        for (int k = 0; k < i; ++k)
            if (s.g[(size_t)k].ordinal == ordinal) {
                set_error("UUID %s listed twice", uuids[i]);
                return GMP_ERR_INVALID;
            }
        s.g[(size_t)i].ordinal = ordinal;
        out->cuda_ordinal[i] = ordinal;
        snprintf(out->uuid[i], GMP_UUID_LEN, "%s", uuids[i]);
    }
    // CUDA's view
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            int can = 1;
            if (i != j) GMP_CUDA(cudaDeviceCanAccessPeer(&can, s.g[(size_t)i].ordinal, s.g[(size_t)j].ordinal));
            out->can_access[cell(n, i, j)] = can;
        }
    // NVML's view; its absence is reported, not fatal (the bandwidth part still runs)
    char keep_err[512];
    if (fill_nvml_matrices(uuids, out) == GMP_OK) {
        out->nvml_available = 1;
        out->matrix_match = 1;
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) {
                if (i == j) continue;
                const int nvml_ok = out->nvml_read[cell(n, i, j)] == 0 && out->nvml_write[cell(n, i, j)] == 0;
                if ((out->can_access[cell(n, i, j)] != 0) != (nvml_ok != 0)) out->matrix_match = 0;
            }
    } else {
        snprintf(keep_err, sizeof(keep_err), "%s", gmp_last_error());
        out->nvml_available = 0;
        out->matrix_match = -1;
        clear_error();
    }
    if (o.flags & GMP_P2P_F_MATRIX_ONLY) {
        out->ns_setup = now_ns() - t0;
        return GMP_OK;
    }
    // Primary contexts first, all at once: one after another they cost ~110 ms each, together the driver overlaps part of
    // the work (8 GPUs: ~0.6 s instead of ~0.9 s; they exist already when the HBM step ran in this process).
    {
        std::vector<std::thread> starters;
        std::vector<cudaError_t> started((size_t)n, cudaSuccess);
        for (int i = 0; i < n; ++i)
            starters.emplace_back([&, i]() {
                cudaError_t e = cudaSetDevice(s.g[(size_t)i].ordinal);
                if (e == cudaSuccess) e = cudaFree(0);
                started[(size_t)i] = e;
            });
        for (std::thread& t : starters) t.join();
        for (int i = 0; i < n; ++i) {
            s.g[(size_t)i].touched = true;  // a context may exist now: teardown must visit the device
            if (started[(size_t)i] != cudaSuccess) return cuda_fail(started[(size_t)i], "primary context");
        }
    }
    // peer mappings, one buffer per GPU
    for (int i = 0; i < n; ++i) {
        PeerGpu& p = s.g[(size_t)i];
        GMP_CUDA(cudaSetDevice(p.ordinal));
        p.touched = true;
        GMP_CUDA(cudaStreamCreateWithFlags(&p.stream, cudaStreamNonBlocking));
        GMP_CUDA(cudaEventCreate(&p.e0));
        GMP_CUDA(cudaEventCreate(&p.e1));
        GMP_CUDA(cudaMalloc(&p.buf, o.bytes));
        GMP_CUDA(cudaMalloc(&p.acc, 4 * sizeof(unsigned long long)));
        for (int j = 0; j < n; ++j) {
            if (i == j || !out->can_access[cell(n, i, j)]) continue;
            cudaError_t e = cudaDeviceEnablePeerAccess(s.g[(size_t)j].ordinal, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
            else if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
            else p.enabled_peers.push_back(s.g[(size_t)j].ordinal);
        }
    }
    out->ns_setup = now_ns() - t0;

    t0 = now_ns();
    int rc = run_pairs(s, o, out);
    out->ns_pairs = now_ns() - t0;
    if (rc != GMP_OK) return rc;
    if (!(o.flags & GMP_P2P_F_SKIP_BIDIR) && n >= 2) {
        t0 = now_ns();
        rc = run_bidir(s, o, out);
        out->ns_bidir = now_ns() - t0;
        if (rc != GMP_OK) return rc;
    }
    if (!(o.flags & GMP_P2P_F_SKIP_ALLTOALL) && n >= 2) {
        t0 = now_ns();
        rc = run_alltoall(s, o, out);
        out->ns_alltoall = now_ns() - t0;
        if (rc != GMP_OK) return rc;
    }
    return GMP_OK;
}

}  // namespace

extern "C" int gmp_probe_p2p(const char* const* uuids, int n, const gmp_p2p_opts* opts, gmp_p2p_result* out) try {
    clear_error();
    if (out == nullptr) {
        set_error("out is NULL");
        return GMP_ERR_INVALID;
    }
    memset(out, 0, sizeof(*out));
    out->first_bad_src = out->first_bad_dst = -1;
    for (int k = 0; k < GMP_P2P_CELLS; ++k) out->atomics_ok[k] = -1;
    for (int i = 0; i < GMP_MAX_DEVICES; ++i) {
        out->cuda_ordinal[i] = -1;
        out->nvlink_active[i] = -1;
    }
    out->matrix_match = -1;
    if (uuids == nullptr || n < 1 || n > GMP_MAX_DEVICES) {
        set_error("need 1..%d UUIDs, got %d", GMP_MAX_DEVICES, n);
        return out->status = GMP_ERR_INVALID;
    }
    for (int i = 0; i < n; ++i)
        if (uuids[i] == nullptr || uuids[i][0] == '\0') {
            set_error("UUID %d is empty", i);
            return out->status = GMP_ERR_INVALID;
        }
    gmp_p2p_opts o;
    memset(&o, 0, sizeof(o));
    if (opts) o = *opts;
    if (o.bytes == 0) o.bytes = GMP_P2P_DEFAULT_BYTES;
    o.bytes &= ~31ull;
    if (o.bytes == 0) {
        set_error("bytes must be at least 32");
        return out->status = GMP_ERR_INVALID;
    }
    if (o.iters == 0) o.iters = 2;
    if (o.seed == 0) o.seed = 0xD1B54A32D192ED03ull;
    out->n = n;
    out->bytes = o.bytes;

    const uint64_t t_begin = now_ns();
    Session s;
    s.g.resize((size_t)n);
    s.keep = (o.flags & GMP_P2P_F_KEEP_DEVICE) != 0;
    (void)cudaGetDevice(&s.prev_device);
    (void)cudaGetLastError();
    int rc = run(uuids, o, out, s);
    char keep[512];
    snprintf(keep, sizeof(keep), "%s", gmp_last_error());
    const uint64_t tt = now_ns();
    s.teardown();
    out->ns_teardown = now_ns() - tt;
    out->ns_total = now_ns() - t_begin;
    if (rc != GMP_OK) {
        set_error("%s", keep);
        return out->status = rc;
    }
    if (out->mismatch_lanes != 0) {
        set_error("%llu mismatching lanes, first transfer GPU %d -> GPU %d", (unsigned long long)out->mismatch_lanes,
                  out->first_bad_src, out->first_bad_dst);
        return out->status = GMP_ERR_VERIFY;
    }
    if (out->matrix_match == 0) {
        set_error("cudaDeviceCanAccessPeer matrix differs from NVML's READ/WRITE P2P status");
        return out->status = GMP_ERR_P2P_MATRIX;
    }
    return out->status = GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

extern "C" int gmp_peer_access(int device, int peer, int enable) try {
    clear_error();
    int prev = -1;
    (void)cudaGetDevice(&prev);
    GMP_CUDA(cudaSetDevice(device));
    cudaError_t e = enable ? cudaDeviceEnablePeerAccess(peer, 0) : cudaDeviceDisablePeerAccess(peer);
    if (e == cudaErrorPeerAccessAlreadyEnabled || e == cudaErrorPeerAccessNotEnabled) {
        (void)cudaGetLastError();
        e = cudaSuccess;
    }
    if (prev >= 0) (void)cudaSetDevice(prev);
    if (e != cudaSuccess) return cuda_fail(e, enable ? "cudaDeviceEnablePeerAccess" : "cudaDeviceDisablePeerAccess");
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}
