// Whole-pod probe (gmp_probe_all) and its JSON line (gmp_report_json): section 5b of
// include/gpumounter_probe.h.  This is what the worker's hook runs, from inside the pod's device
// cgroup and mount namespace, right after the mount loop (ref: pkg/server/gpu-mount/server.go:97-98).

#include "gmp_internal.h"
#include "gmp_nvml_hold.h"

#include <dlfcn.h>
#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <exception>
#include <memory>
#include <new>
#include <string>
#include <thread>
#include <vector>

namespace {

using namespace gmp;

int check_node(const std::string& path, int missing_code, int denied_code) {
    const int fd = open(path.c_str(), O_RDWR | O_CLOEXEC);
    if (fd >= 0) {
        close(fd);
        return GMP_OK;
    }
    const int e = errno;
    set_error("%s: %s", path.c_str(), strerror(e));
    if (e == EPERM || e == EACCES) return denied_code;  // the device cgroup (v1 list or v2 program) or the file mode refuses
    return missing_code;  // ENOENT / ENOTDIR: no node; ENXIO / ENODEV: a node with no device behind it (wrong minor, no driver)
}

int check_nodes(const gmp_device* expected, int n, const char* dev_dir) {
    const std::string dir = (dev_dir && dev_dir[0]) ? dev_dir : "/dev";
    int rc = check_node(dir + "/nvidiactl", GMP_ERR_NODE_CTL, GMP_ERR_NODE_CTL);
    if (rc != GMP_OK) return rc;
    rc = check_node(dir + "/nvidia-uvm", GMP_ERR_NODE_UVM, GMP_ERR_NODE_UVM);
    if (rc != GMP_OK) return rc;
    for (int i = 0; i < n; ++i) {
        if (expected[i].minor == UINT32_MAX) continue;
        rc = check_node(dir + "/nvidia" + std::to_string(expected[i].minor), GMP_ERR_NODE_GPU, GMP_ERR_NODE_DENIED);
        if (rc != GMP_OK) return rc;
    }
    void* h = dlopen("libcuda.so.1", RTLD_LAZY | RTLD_LOCAL);
    if (h == nullptr) {
        set_error("libcuda.so.1: %s", dlerror());
        return GMP_ERR_DRIVER_LIB;
    }
    dlclose(h);
    return GMP_OK;
}

void first_failure(gmp_report* r, int rc) {
    if (rc != GMP_OK && r->status == GMP_OK) {
        r->status = rc;
        snprintf(r->detail, sizeof(r->detail), "%s", gmp_last_error());
    }
}

int find_uuid(const gmp_device* list, int n, const char* uuid) {
    for (int i = 0; i < n; ++i)
        if (strncmp(list[i].uuid, uuid, GMP_UUID_LEN) == 0) return i;
    return -1;
}

// ---- JSON -----------------------------------------------------------------------------------
struct Json {
    std::string s;
    void raw(const char* t) { s += t; }
    void key(const char* k) {
        if (!s.empty() && s.back() != '{' && s.back() != '[') s += ',';
        s += '"';
        s += k;
        s += "\":";
    }
    void sep() {
        if (!s.empty() && s.back() != '{' && s.back() != '[') s += ',';
    }
    void str(const char* v) {
        s += '"';
        for (const char* c = v; *c; ++c) {
            if (*c == '"' || *c == '\\') {
                s += '\\';
                s += *c;
            } else if ((unsigned char)*c < 0x20) {
                char esc[8];
                snprintf(esc, sizeof(esc), "\\u%04x", (unsigned char)*c);
                s += esc;
            } else if ((unsigned char)*c < 0x80) {
                s += *c;
            } else {
                // a well-formed UTF-8 sequence passes through; anything else (a path or message in another encoding)
                // becomes U+FFFD so that the line is always valid JSON
                const unsigned char b0 = (unsigned char)*c;
                const int len = (b0 & 0xE0) == 0xC0 ? 2 : (b0 & 0xF0) == 0xE0 ? 3 : (b0 & 0xF8) == 0xF0 ? 4 : 0;
                bool ok = len != 0 && b0 >= 0xC2 && b0 <= 0xF4;
                for (int k = 1; ok && k < len; ++k) ok = ((unsigned char)c[k] & 0xC0) == 0x80;
                if (ok && len == 3) ok = !(b0 == 0xE0 && (unsigned char)c[1] < 0xA0) && !(b0 == 0xED && (unsigned char)c[1] >= 0xA0);
                if (ok && len == 4) ok = !(b0 == 0xF0 && (unsigned char)c[1] < 0x90) && !(b0 == 0xF4 && (unsigned char)c[1] >= 0x90);
                if (ok) {
                    s.append(c, (size_t)len);
                    c += len - 1;
                } else {
                    s += "\\ufffd";
                }
            }
        }
        s += '"';
    }
    void num(double v) {
        if (!(v == v) || v > 1e300 || v < -1e300) {  // NaN / infinity have no JSON spelling
            s += "null";
            return;
        }
        char b[360];
        snprintf(b, sizeof(b), "%.3f", v);
        s += b;
    }
    void i64(long long v) { s += std::to_string(v); }
    void u64(unsigned long long v) { s += std::to_string(v); }
    void kstr(const char* k, const char* v) { key(k); str(v); }
    void knum(const char* k, double v) { key(k); num(v); }
    void ki(const char* k, long long v) { key(k); i64(v); }
    void ku(const char* k, unsigned long long v) { key(k); u64(v); }
    void kms(const char* k, uint64_t ns) { key(k); num((double)ns / 1e6); }
    template <typename T>
    void matrix(const char* k, const T* m, int n, bool is_float) {
        key(k);
        s += '[';
        for (int i = 0; i < n; ++i) {
            if (i) s += ',';
            s += '[';
            for (int j = 0; j < n; ++j) {
                if (j) s += ',';
                if (is_float) num((double)m[i * n + j]);
                else i64((long long)m[i * n + j]);
            }
            s += ']';
        }
        s += ']';
    }
};

}  // namespace

namespace {
int probe_all_impl(const gmp_device* expected, int n_expected, const gmp_probe_opts* opts, gmp_report* out);
}

extern "C" int gmp_probe_all(const gmp_device* expected, int n_expected, const gmp_probe_opts* opts, gmp_report* out) {
    try {  // std::thread / std::string / std::vector may throw; nothing crosses the C ABI
        return probe_all_impl(expected, n_expected, opts, out);
    } catch (const std::bad_alloc&) {
        set_error("out of memory");
        if (out) out->status = GMP_ERR_NOMEM;
        return GMP_ERR_NOMEM;
    } catch (const std::exception& e) {
        set_error("gmp_probe_all: %s", e.what());
        if (out) out->status = GMP_ERR_SYS;
        return GMP_ERR_SYS;
    }
}

namespace {
int probe_all_impl(const gmp_device* expected, int n_expected, const gmp_probe_opts* opts, gmp_report* out) {
    clear_error();
    if (out == nullptr || n_expected < 0 || n_expected > GMP_MAX_DEVICES || (n_expected > 0 && expected == nullptr)) {
        set_error("bad argument");
        return GMP_ERR_INVALID;
    }
    memset(out, 0, sizeof(*out));
    gmp_probe_opts o;
    memset(&o, 0, sizeof(o));
    if (opts) o = *opts;
    if (opts == nullptr) o.hbm.dma_bytes = UINT64_MAX;
    const uint64_t t_begin = now_ns();
    out->mono_start_ns = t_begin;

    // 1. device nodes and the driver library, as seen from this process's namespaces
    uint64_t t0 = now_ns();
    if (o.n_also_visible < 0 || o.n_also_visible > GMP_MAX_DEVICES || (o.n_also_visible > 0 && o.also_visible == nullptr)) {
        set_error("bad also_visible list");
        return GMP_ERR_INVALID;
    }
    if (!(o.flags & GMP_PROBE_F_SKIP_NODES)) {
        out->nodes_status = check_nodes(expected, n_expected, o.dev_dir);
        if (out->nodes_status == GMP_OK && o.n_also_visible > 0) {
            // the GPUs the pod held before this mount: their nodes must still open as well (control nodes checked above)
            const std::string dir = (o.dev_dir && o.dev_dir[0]) ? o.dev_dir : "/dev";
            for (int i = 0; i < o.n_also_visible && out->nodes_status == GMP_OK; ++i)
                if (o.also_visible[i].minor != UINT32_MAX)
                    out->nodes_status = check_node(dir + "/nvidia" + std::to_string(o.also_visible[i].minor), GMP_ERR_NODE_GPU,
                                                   GMP_ERR_NODE_DENIED);
        }
        first_failure(out, out->nodes_status);
    }
    out->ns_nodes = now_ns() - t0;
    if (out->status != GMP_OK || (o.flags & GMP_PROBE_F_NODES_ONLY)) {
        if (out->status == GMP_OK) {
            out->n = n_expected;
            for (int i = 0; i < n_expected; ++i) out->devices[i] = expected[i];
            out->nvml_status = GMP_ERR_UNAVAILABLE;  // not consulted
        }
        out->ns_total = now_ns() - t_begin;
        return out->status;
    }

    // 2. device set.  NVML is loaded and queried on its own thread so that its ~100 ms start-up overlaps cuInit;
    //    its answer is only needed for the (UUID, minor) cross-check after the kernels have run.
    t0 = now_ns();
    gmp_device nvml[GMP_MAX_DEVICES];
    int n_nvml = 0;
    std::string nvml_error;
    // One NVML attach for the whole probe: the enumeration, the per-GPU health samples and the P2P status matrices each
    // bracket themselves with nvmlInit_v2/nvmlShutdown, and a cold attach costs 30-140 ms (8-GPU host: 140 ms, i.e. over a
    // second for 8 health samples).  Taken on the NVML thread so that it overlaps cuInit; released when the probe returns.
    std::unique_ptr<NvmlHold> nvml_hold;
    std::thread nvml_thread;
    if (o.flags & GMP_PROBE_F_SKIP_NVML) {
        out->nvml_status = GMP_ERR_UNAVAILABLE;
    } else {
        nvml_thread = std::thread([&]() {
            const uint64_t tn = now_ns();
            nvml_hold.reset(new (std::nothrow) NvmlHold());
            out->nvml_status = gmp_enumerate_nvml(nvml, GMP_MAX_DEVICES, &n_nvml);
            if (out->nvml_status == GMP_ERR_NVML && n_expected > 0) {
                // Inside a pod that holds only some of the node's GPUs NVML's index walk stops at the first GPU the
                // pod's device cgroup hides; the mounted ones still answer by UUID, which is all the cross-check needs.
                for (int i = 0; i < n_expected; ++i) {
                    nvml[i] = expected[i];
                    nvml[i].cuda_ordinal = -1;
                }
                int found = 0;
                if (nvml_lookup_by_uuid(nvml, n_expected, &found) == GMP_OK) {
                    n_nvml = n_expected;
                    out->nvml_status = GMP_OK;
                }
            }
            if (out->nvml_status != GMP_OK) {
                n_nvml = 0;  // reported, not fatal: a pod need not ship libnvidia-ml
                nvml_error = gmp_last_error();
            }
            out->ns_nvml = now_ns() - tn;
        });
    }
    struct Joiner {
        std::thread& t;
        ~Joiner() {
            if (t.joinable()) t.join();
        }
    } joiner{nvml_thread};

    gmp_device cuda[GMP_MAX_DEVICES];
    int n_cuda = 0;
    int rc = gmp_enumerate_cuda(cuda, GMP_MAX_DEVICES, &n_cuda);
    out->ns_cuda_init = now_ns() - t0;
    if (rc != GMP_OK) {
        out->device_set_status = rc;
        first_failure(out, rc);
        out->ns_total = now_ns() - t_begin;
        return out->status;
    }
    if (n_expected > 0) {
        // the set to probe is the expected one; CUDA must see exactly these UUIDs
        out->n = n_expected;
        for (int i = 0; i < n_expected; ++i) {
            out->devices[i] = expected[i];
            out->devices[i].cuda_ordinal = -1;
            const int c = find_uuid(cuda, n_cuda, expected[i].uuid);
            if (c < 0) {
                set_error("expected %s is not visible to CUDA (%d visible)", expected[i].uuid, n_cuda);
                out->device_set_status = GMP_ERR_DEVICE_SET;
            } else {
                out->devices[i].cuda_ordinal = cuda[c].cuda_ordinal;
                if (out->devices[i].pci_bus_id[0] == '\0') memcpy(out->devices[i].pci_bus_id, cuda[c].pci_bus_id, GMP_PCI_LEN);
            }
        }
        for (int i = 0; i < o.n_also_visible && out->device_set_status == GMP_OK; ++i)
            if (find_uuid(cuda, n_cuda, o.also_visible[i].uuid) < 0) {
                set_error("%s, which the pod held before this mount, is not visible to CUDA (%d visible)", o.also_visible[i].uuid, n_cuda);
                out->device_set_status = GMP_ERR_DEVICE_SET;
            }
        if (out->device_set_status == GMP_OK && n_cuda != n_expected + o.n_also_visible && !(o.flags & GMP_PROBE_F_HOST_VIEW)) {
            set_error("CUDA sees %d devices, %d were mounted (%d now + %d held before)", n_cuda, n_expected + o.n_also_visible,
                      n_expected, o.n_also_visible);
            out->device_set_status = GMP_ERR_DEVICE_SET;
        }
    } else {
        out->n = n_cuda;
        for (int i = 0; i < n_cuda; ++i) out->devices[i] = cuda[i];
    }
    first_failure(out, out->device_set_status);
    out->ns_enumerate = now_ns() - t0;
    if (out->status != GMP_OK) {
        out->ns_total = now_ns() - t_begin;
        return out->status;
    }

    // 3. HBM probe, one host thread per GPU (independent units, no communication)
    const int n = out->n;
    t0 = now_ns();
    if (!(o.flags & GMP_PROBE_F_SKIP_HBM)) {
        gmp_hbm_opts h = o.hbm;
        h.flags |= GMP_HBM_F_KEEP_DEVICE;  // contexts are needed again for the P2P step; reset happens at the end
        std::vector<std::thread> workers;
        std::vector<std::string> errors((size_t)n);
        for (int i = 0; i < n; ++i) {
            workers.emplace_back([&, i]() {
                gmp_probe_hbm(out->devices[i].uuid, &h, &out->hbm[i]);
                errors[(size_t)i] = gmp_last_error();
            });
        }
        for (std::thread& w : workers) w.join();
        for (int i = 0; i < n; ++i) {
            if (out->hbm[i].status != GMP_OK && out->status == GMP_OK) {
                out->status = out->hbm[i].status;
                snprintf(out->detail, sizeof(out->detail), "%s: %s", out->devices[i].uuid, errors[(size_t)i].c_str());
            }
        }
    }
    out->ns_hbm = now_ns() - t0;
    for (int i = 0; i < n; ++i)
        if (out->hbm[i].mono_first_verified_ns > out->mono_usable_ns) out->mono_usable_ns = out->hbm[i].mono_first_verified_ns;
    for (int i = 0; i < n; ++i)
        if (!(o.flags & GMP_PROBE_F_SKIP_HBM) && out->hbm[i].mono_first_verified_ns == 0) out->mono_usable_ns = 0;

    // 4. NVML's view joins in: minor numbers and NVML indices by UUID, and the (UUID, minor) cross-check
    if (nvml_thread.joinable()) nvml_thread.join();
    for (int i = 0; i < n; ++i) {  // clocks / throttle reasons / ECC while the GPUs are still warm from the passes
        memset(&out->health[i], 0xFF, sizeof(out->health[i]));
        if (n_nvml > 0) (void)gmp_nvml_health(out->devices[i].uuid, &out->health[i]);
    }
    clear_error();
    for (int i = 0; i < n && n_nvml > 0; ++i) {
        const int k = find_uuid(nvml, n_nvml, out->devices[i].uuid);
        if (k < 0) {
            if (out->device_set_status == GMP_OK) {
                set_error("%s is not in NVML's enumeration", out->devices[i].uuid);
                out->device_set_status = GMP_ERR_DEVICE_SET;
            }
            continue;
        }
        if (out->devices[i].minor != UINT32_MAX && out->devices[i].minor != nvml[k].minor && out->device_set_status == GMP_OK) {
            set_error("%s: mounted as minor %u, NVML says %u", out->devices[i].uuid, out->devices[i].minor, nvml[k].minor);
            out->device_set_status = GMP_ERR_DEVICE_SET;
        }
        if (out->devices[i].minor == UINT32_MAX) out->devices[i].minor = nvml[k].minor;
        out->devices[i].nvml_index = nvml[k].nvml_index;
    }
    first_failure(out, out->device_set_status);

    // 5. P2P across the set
    t0 = now_ns();
    const bool keep = (o.flags & GMP_PROBE_F_KEEP_DEVICE) != 0;
    bool devices_reset = false;
    if (!(o.flags & GMP_PROBE_F_SKIP_P2P) && n >= 2 && out->status == GMP_OK) {
        const char* uuids[GMP_MAX_DEVICES];
        for (int i = 0; i < n; ++i) uuids[i] = out->devices[i].uuid;
        gmp_p2p_opts p = o.p2p;
        if (keep) p.flags |= GMP_P2P_F_KEEP_DEVICE;
        if (o.flags & GMP_PROBE_F_SKIP_NVML) {
            // matrices cannot be cross-checked without NVML; bandwidth and verification still run
        }
        rc = gmp_probe_p2p(uuids, n, &p, &out->p2p);
        first_failure(out, rc);
        devices_reset = !keep;
    }
    out->ns_p2p = now_ns() - t0;
    if (!keep && !devices_reset && !(o.flags & GMP_PROBE_F_SKIP_HBM)) {
        int prev = -1;
        (void)cudaGetDevice(&prev);
        for (int i = 0; i < n; ++i) {
            if (out->devices[i].cuda_ordinal < 0) continue;
            if (cudaSetDevice(out->devices[i].cuda_ordinal) == cudaSuccess) {
                (void)cudaDeviceReset();
                forget_device_state(out->devices[i].cuda_ordinal);
            }
        }
        if (prev >= 0) (void)cudaSetDevice(prev);
        (void)cudaGetLastError();
    }
    out->ns_total = now_ns() - t_begin;
    return out->status;
}
}  // namespace

extern "C" int gmp_verify_gone(const gmp_device* gone, int n, const char* dev_dir) try {
    clear_error();
    if (n < 0 || (n > 0 && gone == nullptr)) {
        set_error("bad argument");
        return GMP_ERR_INVALID;
    }
    const std::string dir = (dev_dir && dev_dir[0]) ? dev_dir : "/dev";
    for (int i = 0; i < n; ++i) {
        if (gone[i].minor == UINT32_MAX) {
            set_error("%s: minor number not given", gone[i].uuid);
            return GMP_ERR_INVALID;
        }
        const std::string path = dir + "/nvidia" + std::to_string(gone[i].minor);
        const int fd = open(path.c_str(), O_RDWR | O_CLOEXEC);
        if (fd >= 0) {
            close(fd);
            set_error("%s (%s) still exists and still opens", path.c_str(), gone[i].uuid);
            return GMP_ERR_STILL_MOUNTED;
        }
        if (errno != ENOENT && errno != ENOTDIR) {
            set_error("%s (%s) still exists (open: %s)", path.c_str(), gone[i].uuid, strerror(errno));
            return GMP_ERR_STILL_MOUNTED;
        }
    }
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

extern "C" int gmp_report_apply_thresholds(gmp_report* r, double min_hbm_gbs, double min_p2p_gbs) {
    clear_error();
    if (r == nullptr) {
        set_error("report is NULL");
        return GMP_ERR_INVALID;
    }
    if (r->status != GMP_OK) return r->status;
    auto fail = [r](const char* what, const char* uuid, int src, int dst, double got, double want) {
        if (uuid) snprintf(r->detail, sizeof(r->detail), "%s: %s %.1f GB/s is under the threshold of %.1f", uuid, what, got, want);
        else snprintf(r->detail, sizeof(r->detail), "GPU %d -> GPU %d: peer %s %.1f GB/s is under the threshold of %.1f", src, dst, what, got, want);
        set_error("%s", r->detail);
        return r->status = GMP_ERR_SLOW;
    };
    if (min_hbm_gbs > 0.0) {
        for (int i = 0; i < r->n && i < GMP_MAX_DEVICES; ++i) {
            const gmp_hbm_result& h = r->hbm[i];
            if (h.write_gbs_best > 0.0 && h.write_gbs_best < min_hbm_gbs) return fail("write", h.uuid, 0, 0, h.write_gbs_best, min_hbm_gbs);
            if (h.read_gbs_best > 0.0 && h.read_gbs_best < min_hbm_gbs) return fail("read-verify", h.uuid, 0, 0, h.read_gbs_best, min_hbm_gbs);
            if (h.copy_gbs_best > 0.0 && h.copy_gbs_best < min_hbm_gbs) return fail("copy", h.uuid, 0, 0, h.copy_gbs_best, min_hbm_gbs);
        }
    }
    if (min_p2p_gbs > 0.0 && r->p2p.n >= 2 && r->p2p.n <= GMP_MAX_DEVICES) {
        const int n = r->p2p.n;
        for (int a = 0; a < n; ++a)
            for (int b = 0; b < n; ++b) {
                if (a == b) continue;
                const double w = r->p2p.write_gbs[a * n + b], l = r->p2p.read_gbs[a * n + b];
                if (w > 0.0 && w < min_p2p_gbs) return fail("store", nullptr, a, b, w, min_p2p_gbs);
                if (l > 0.0 && l < min_p2p_gbs) return fail("load", nullptr, a, b, l, min_p2p_gbs);
            }
    }
    return r->status;
}

extern "C" int gmp_report_check_health(gmp_report* r) {
    clear_error();
    if (r == nullptr) {
        set_error("report is NULL");
        return GMP_ERR_INVALID;
    }
    if (r->status != GMP_OK) return r->status;
    for (int i = 0; i < r->n && i < GMP_MAX_DEVICES; ++i) {
        const gmp_gpu_health& h = r->health[i];
        const char* why = nullptr;
        char buf[96];
        if (h.ecc_uncorrected_volatile != UINT64_MAX && h.ecc_uncorrected_volatile > 0) {
            snprintf(buf, sizeof(buf), "%llu uncorrected ECC errors since the last driver load", (unsigned long long)h.ecc_uncorrected_volatile);
            why = buf;
        } else if (h.remap_failure != UINT32_MAX && h.remap_failure != 0) {
            why = "HBM row remapping has failed (no spare rows left)";
        } else if (h.remap_pending != UINT32_MAX && h.remap_pending != 0) {
            why = "an HBM row remap is pending: the GPU needs a reset";
        } else if (h.throttle_reasons != UINT64_MAX && (h.throttle_reasons & 0x8ull)) {
            why = "hardware slowdown is active";
        } else if (h.throttle_reasons != UINT64_MAX && (h.throttle_reasons & 0x40ull)) {
            why = "hardware thermal slowdown is active";
        } else if (h.throttle_reasons != UINT64_MAX && (h.throttle_reasons & 0x80ull)) {
            why = "hardware power brake is active";
        }
        if (why != nullptr) {
            snprintf(r->detail, sizeof(r->detail), "%s: %s", r->devices[i].uuid, why);
            set_error("%s", r->detail);
            return r->status = GMP_ERR_UNHEALTHY;
        }
    }
    return r->status;
}

extern "C" int gmp_report_json(const gmp_report* r, char* buf, size_t cap) try {
    clear_error();
    if (r == nullptr || buf == nullptr) {
        set_error("NULL argument");
        return GMP_ERR_INVALID;
    }
    Json j;
    const int n = r->n;
    j.raw("{");
    j.ki("status", r->status);
    j.kstr("verdict", gmp_strerror(r->status));
    j.kstr("detail", r->detail);
    j.ki("n_gpus", n);
    j.ki("nodes_status", r->nodes_status);
    j.ki("device_set_status", r->device_set_status);
    j.ki("nvml_status", r->nvml_status);
    j.key("devices");
    j.raw("[");
    for (int i = 0; i < n; ++i) {
        j.sep();
        j.raw("{");
        j.kstr("uuid", r->devices[i].uuid);
        j.ki("minor", r->devices[i].minor == UINT32_MAX ? -1 : (long long)r->devices[i].minor);
        j.kstr("pci_bus_id", r->devices[i].pci_bus_id);
        j.ki("nvml_index", r->devices[i].nvml_index);
        j.ki("cuda_ordinal", r->devices[i].cuda_ordinal);
        j.raw("}");
    }
    j.raw("]");
    j.key("hbm");
    j.raw("[");
    for (int i = 0; i < n; ++i) {
        const gmp_hbm_result& h = r->hbm[i];
        if (h.uuid[0] == '\0' && h.status == GMP_OK) continue;  // step skipped (SKIP_HBM, NODES_ONLY): no entry
        j.sep();
        j.raw("{");
        j.kstr("uuid", h.uuid);
        j.ki("status", h.status);
        j.ku("bytes", h.bytes);
        j.knum("write_gbs", h.write_gbs_best);
        j.knum("read_gbs", h.read_gbs_best);
        j.knum("copy_gbs", h.copy_gbs_best);
        j.knum("write_gbs_median", h.write_gbs_median);
        j.knum("read_gbs_median", h.read_gbs_median);
        j.knum("copy_gbs_median", h.copy_gbs_median);
        j.knum("h2d_gbs", h.h2d_gbs);
        j.knum("d2h_gbs", h.d2h_gbs);
        j.knum("lib_memset_gbs", h.lib_memset_gbs);
        j.knum("lib_memcpy_gbs", h.lib_memcpy_gbs);
        j.ku("mismatch_lanes", h.mismatch_lanes);
        j.ku("xor64", h.xor64);
        j.ku("sum64", h.sum64);
        j.kms("ms_select_device", h.ns_select_device);
        j.kms("ms_context", h.ns_context);
        j.kms("ms_alloc", h.ns_alloc);
        j.kms("ms_first_kernel", h.ns_first_kernel);
        j.kms("ms_kernels", h.ns_kernels);
        j.kms("ms_dma", h.ns_dma);
        j.kms("ms_teardown", h.ns_teardown);
        j.kms("ms_total", h.ns_total);
        j.ku("mono_first_verified_ns", h.mono_first_verified_ns);
        j.ku("mono_context_begin_ns", h.mono_context_begin_ns);
        j.ku("mono_context_end_ns", h.mono_context_end_ns);
        j.raw("}");
    }
    j.raw("]");
    j.key("health");
    j.raw("[");
    for (int i = 0; i < n; ++i) {
        const gmp_gpu_health& h = r->health[i];
        static const gmp_gpu_health never_sampled = {};  // the report starts zeroed; step 4 fills 0xFF or real values
        if (memcmp(&h, &never_sampled, sizeof(h)) == 0) continue;  // NODES_ONLY, or a failed node check
        auto opt32 = [](uint32_t v) { return v == UINT32_MAX ? -1ll : (long long)v; };
        auto opt64 = [](uint64_t v) { return v == UINT64_MAX ? -1ll : (long long)v; };
        j.sep();
        j.raw("{");
        j.ki("sm_mhz", opt32(h.sm_mhz));
        j.ki("mem_mhz", opt32(h.mem_mhz));
        j.ki("sm_max_mhz", opt32(h.sm_max_mhz));
        j.ki("mem_max_mhz", opt32(h.mem_max_mhz));
        j.ki("temperature_c", opt32(h.temperature_c));
        j.ki("power_mw", opt32(h.power_mw));
        j.ki("throttle_reasons", opt64(h.throttle_reasons));
        j.ki("ecc_uncorrected_volatile", opt64(h.ecc_uncorrected_volatile));
        j.ki("ecc_corrected_volatile", opt64(h.ecc_corrected_volatile));
        j.ki("remapped_rows_correctable", opt32(h.remapped_rows_correctable));
        j.ki("remapped_rows_uncorrectable", opt32(h.remapped_rows_uncorrectable));
        j.ki("remap_pending", opt32(h.remap_pending));
        j.ki("remap_failure", opt32(h.remap_failure));
        j.ki("nvlink_crc_flit_errors", opt64(h.nvlink_crc_flit_errors));
        j.ki("nvlink_crc_data_errors", opt64(h.nvlink_crc_data_errors));
        j.ki("nvlink_replay_errors", opt64(h.nvlink_replay_errors));
        j.ki("nvlink_recovery_errors", opt64(h.nvlink_recovery_errors));
        j.raw("}");
    }
    j.raw("]");
    if (r->p2p.n >= 2) {
        const gmp_p2p_result& p = r->p2p;
        j.key("p2p");
        j.raw("{");
        j.ki("status", p.status);
        j.ki("matrix_match", p.matrix_match);
        j.ki("nvml_available", p.nvml_available);
        j.ku("bytes", p.bytes);
        j.ku("mismatch_lanes", p.mismatch_lanes);
        j.matrix("can_access", p.can_access, p.n, false);
        j.matrix("nvml_read", p.nvml_read, p.n, false);
        j.matrix("nvml_write", p.nvml_write, p.n, false);
        j.matrix("nvml_nvlink", p.nvml_nvlink, p.n, false);
        j.matrix("nvml_atomics", p.nvml_atomics, p.n, false);
        j.matrix("write_gbs", p.write_gbs, p.n, true);
        j.matrix("read_gbs", p.read_gbs, p.n, true);
        j.matrix("memcpy_peer_gbs", p.memcpy_gbs, p.n, true);
        j.matrix("atomics_ok", p.atomics_ok, p.n, false);
        j.key("bidir_gbs");
        j.raw("[");
        for (int i = 0; i < p.n; ++i) {
            if (i) j.raw(",");
            j.num(p.bidir_gbs[i]);
        }
        j.raw("]");
        j.key("nvlink_active");
        j.raw("[");
        for (int i = 0; i < p.n; ++i) {
            if (i) j.raw(",");
            j.i64(p.nvlink_active[i]);
        }
        j.raw("]");
        j.knum("alltoall_gbs", p.alltoall_gbs);
        // NVSwitch gives every pair the same bandwidth: a pair well below the others is the finding an operator wants
        // named (a degraded link, a GPU routed over PCIe).  Per kind, pairs under 80 % of that kind's median.
        j.key("slow_pairs");
        j.raw("[");
        {
            const struct {
                const char* kind;
                const double* m;
            } kinds[] = {{"store", p.write_gbs}, {"load", p.read_gbs}, {"memcpy_peer", p.memcpy_gbs}};
            for (const auto& k : kinds) {
                std::vector<double> measured;
                for (int a = 0; a < p.n; ++a)
                    for (int b = 0; b < p.n; ++b)
                        if (a != b && k.m[a * p.n + b] > 0.0) measured.push_back(k.m[a * p.n + b]);
                if (measured.size() < 2) continue;
                std::sort(measured.begin(), measured.end());
                const double median = measured[measured.size() / 2];
                for (int a = 0; a < p.n; ++a)
                    for (int b = 0; b < p.n; ++b) {
                        const double v = k.m[a * p.n + b];
                        if (a == b || !(v > 0.0) || v >= 0.8 * median) continue;
                        j.sep();
                        j.raw("{");
                        j.ki("src", a);
                        j.ki("dst", b);
                        j.kstr("kind", k.kind);
                        j.knum("gbs", v);
                        j.knum("median_gbs", median);
                        j.raw("}");
                    }
            }
        }
        j.raw("]");
        j.kms("ms_setup", p.ns_setup);
        j.kms("ms_pairs", p.ns_pairs);
        j.kms("ms_bidir", p.ns_bidir);
        j.kms("ms_alltoall", p.ns_alltoall);
        j.kms("ms_teardown", p.ns_teardown);
        j.kms("ms_total", p.ns_total);
        j.raw("}");
    }
    j.kms("ms_nodes", r->ns_nodes);
    j.kms("ms_enumerate", r->ns_enumerate);
    j.kms("ms_hbm", r->ns_hbm);
    j.kms("ms_p2p", r->ns_p2p);
    j.kms("ms_cuda_init", r->ns_cuda_init);
    j.kms("ms_nvml", r->ns_nvml);
    j.kms("ms_to_usable", r->mono_usable_ns > r->mono_start_ns ? r->mono_usable_ns - r->mono_start_ns : 0);
    j.ku("mono_start_ns", r->mono_start_ns);
    j.ku("mono_usable_ns", r->mono_usable_ns);
    j.kms("ms_total", r->ns_total);
    j.raw("}");
    if (j.s.size() + 1 > cap) {
        set_error("JSON needs %zu bytes, capacity %zu", j.s.size() + 1, cap);
        return GMP_ERR_BUFFER_TOO_SMALL;
    }
    memcpy(buf, j.s.c_str(), j.s.size() + 1);
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}
