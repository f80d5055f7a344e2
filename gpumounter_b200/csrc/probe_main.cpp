// gpumounter-probe — the standalone form of gmp_probe_all().
//
// The worker cannot prove a mount from its own context (it is privileged and sees every GPU,
// ref: deploy/gpu-mounter-workers.yaml:23-24, docker/GPUMounter-worker/Dockerfile:17), so after
// util.MountGPU it runs THIS binary inside the target: join the pod's device cgroup, setns() into
// its mount namespace, fexecve() the probe from a descriptor opened beforehand.  The binary is
// therefore self-contained: CUDA runtime, libstdc++ and libgcc are linked statically; libcuda.so.1
// and libnvidia-ml.so.1 are resolved from the pod's own root at run time.
//
// Output: one JSON line on stdout.  Exit status: 0 = usable, otherwise -(GMP_ERR_* code).

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "gpumounter_probe.h"

// ---- keeping the glibc floor low ----------------------------------------------------------------------
// The binary is fexecve'd inside arbitrary pods, so it must start against THEIR libc.  Built here against glibc 2.39
// it would otherwise ask for GLIBC_2.38 (strtoul & co. are redirected to __isoc23_* by the headers), GLIBC_2.36
// (arc4random, referenced by the static libstdc++'s std::random_device) and GLIBC_2.35 (_dl_find_object, used by the
// static libgcc unwinder as a fast path).  Own parsers and two local definitions bring the requirement down to
// GLIBC_2.34 (the release that merged libpthread/libdl into libc: unavoidable without an older sysroot).
extern "C" {
unsigned int arc4random(void) {  // never reached by this program (no std::random_device); must merely resolve locally
    unsigned int v = 0x9E3779B9u;
    FILE* f = fopen("/dev/urandom", "rb");
    if (f) {
        if (fread(&v, sizeof(v), 1, f) != 1) v ^= 0x85EBCA6Bu;
        fclose(f);
    }
    return v;
}
int _dl_find_object(void*, void*) { return -1; }  // "not found": the unwinder falls back to dl_iterate_phdr
// the static libstdc++ (eh_alloc.o, parsing GLIBCXX_TUNABLES) was itself compiled against the C23 redirect
unsigned long __isoc23_strtoul(const char* s, char** end, int base) {
    while (*s == ' ' || *s == '\t') ++s;
    if (*s == '+') ++s;
    if ((base == 0 || base == 16) && s[0] == '0' && (s[1] == 'x' || s[1] == 'X')) {
        s += 2;
        base = 16;
    } else if (base == 0) {
        base = s[0] == '0' ? 8 : 10;
    }
    unsigned long v = 0;
    for (;; ++s) {
        int d;
        if (*s >= '0' && *s <= '9') d = *s - '0';
        else if (*s >= 'a' && *s <= 'z') d = *s - 'a' + 10;
        else if (*s >= 'A' && *s <= 'Z') d = *s - 'A' + 10;
        else break;
        if (d >= base) break;
        v = v * (unsigned long)base + (unsigned long)d;
    }
    if (end) *end = const_cast<char*>(s);
    return v;
}
}

namespace {

// unsigned integer in decimal, 0x-hex or 0-octal; *end receives the first unparsed character
unsigned long long parse_u64(const char* s, const char** end) {
    unsigned long long v = 0;
    int base = 10;
    if (s[0] == '0' && (s[1] == 'x' || s[1] == 'X')) {
        base = 16;
        s += 2;
    } else if (s[0] == '0' && s[1] >= '0' && s[1] <= '7') {
        base = 8;
        s += 1;
    }
    for (;; ++s) {
        int d;
        if (*s >= '0' && *s <= '9') d = *s - '0';
        else if (base == 16 && *s >= 'a' && *s <= 'f') d = *s - 'a' + 10;
        else if (base == 16 && *s >= 'A' && *s <= 'F') d = *s - 'A' + 10;
        else break;
        if (d >= base) break;
        v = v * (unsigned)base + (unsigned)d;
    }
    if (end) *end = s;
    return v;
}

// "<number>[.<fraction>][K|M|G]"
unsigned long long parse_size(const char* s) {
    const char* end = nullptr;
    const unsigned long long whole = parse_u64(s, &end);
    double v = (double)whole;
    if (end == s) {
        fprintf(stderr, "bad size '%s'\n", s);
        exit(64);
    }
    if (*end == '.') {
        double scale = 0.1;
        for (++end; *end >= '0' && *end <= '9'; ++end, scale *= 0.1) v += (*end - '0') * scale;
    }
    unsigned long long mult = 1;
    if (*end) {
        switch (*end) {
            case 'k': case 'K': mult = 1ull << 10; break;
            case 'm': case 'M': mult = 1ull << 20; break;
            case 'g': case 'G': mult = 1ull << 30; break;
            default: fprintf(stderr, "bad size '%s'\n", s); exit(64);
        }
        if (end[1] != '\0') {
            fprintf(stderr, "bad size '%s'\n", s);
            exit(64);
        }
    }
    return (unsigned long long)(v * (double)mult);
}

unsigned parse_u32(const char* s) { return (unsigned)parse_u64(s, nullptr); }

void add_device(std::vector<gmp_device>* list, const std::string& spec) {
    gmp_device d;
    memset(&d, 0, sizeof(d));
    d.minor = UINT32_MAX;
    d.nvml_index = d.cuda_ordinal = -1;
    const size_t colon = spec.rfind(':');
    std::string uuid = spec;
    if (colon != std::string::npos && colon + 1 < spec.size() && spec.find_first_not_of("0123456789", colon + 1) == std::string::npos) {
        uuid = spec.substr(0, colon);
        d.minor = parse_u32(spec.c_str() + colon + 1);
    }
    snprintf(d.uuid, sizeof(d.uuid), "%s", uuid.c_str());
    list->push_back(d);
}

void usage() {
    fprintf(stderr,
            "usage: gpumounter-probe --gone GPU-...:minor [--gone ...] [--dev-dir DIR]     (after an unmount)\n"
            "       gpumounter-probe [--uuid GPU-...[:minor]]... [--uuids a,b,c] [--bytes N[K|M|G]] [--iters N]\n"
            "         [--also-mounted GPU-...[:minor]]...  (held by the pod already: must be visible too, not probed)\n"
            "         [--patterns MASK|all] [--dma-bytes N] [--p2p-bytes N] [--p2p-iters N] [--dev-dir DIR]\n"
            "         [--skip-nodes] [--skip-hbm] [--skip-p2p] [--skip-nvml] [--skip-copy] [--matrix-only] [--lib-baselines]\n"
            "         [--min-hbm-gbs N] [--min-p2p-gbs N]  (fail with -25 when a measured rate is lower)\n"
            "         [--strict-health]  (fail with -26 on uncorrected ECC errors or a hardware slowdown)\n"
            "         [--nodes-only]  (device nodes + driver library only: no CUDA, a few ms)\n"
            "         [--host-view]   (run from the host: CUDA may see more GPUs than the listed ones)\n"
            "         [--tune knob=value]...  (launch geometry, see gmp_tune)\n"
            "Probes the GPUs visible to this process (or exactly the listed ones) and prints one JSON line.\n"
            "Defaults may come from GPUMOUNTER_PROBE_BYTES / _ITERS / _PATTERNS / _SEED / _DMA_BYTES / _P2P_BYTES / _SKIP_P2P /\n"
            "_MIN_HBM_GBS / _MIN_P2P_GBS.\n");
}

}  // namespace

int main(int argc, char** argv) {
    // The probe drives one stream per GPU.  The driver sizes a context's channel pool by these two variables and
    // context creation time follows it (B200, driver 580: 8 connections (default) ~190-240 ms, 1 connection
    // ~110-140 ms, 32 connections ~1000 ms per context).  Contexts of one process are created one after another
    // inside the driver, so with eight GPUs this is 0.5 s of mount-to-usable latency.  An operator's own setting wins.
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "1", 0);
    setenv("CUDA_DEVICE_MAX_COPY_CONNECTIONS", "1", 0);
    gmp_probe_opts o;
    memset(&o, 0, sizeof(o));
    o.hbm.dma_bytes = UINT64_MAX;  // library default
    o.flags |= GMP_PROBE_F_KEEP_DEVICE;  // a process of its own: exit releases everything, no explicit reset needed
    double min_hbm_gbs = 0.0, min_p2p_gbs = 0.0;  // acceptance thresholds; 0 = report only
    bool strict_health = false;
    if (const char* v = getenv("GPUMOUNTER_PROBE_STRICT_HEALTH")) strict_health = parse_u32(v) != 0;
    if (const char* v = getenv("GPUMOUNTER_PROBE_MIN_HBM_GBS")) min_hbm_gbs = (double)parse_u32(v);
    if (const char* v = getenv("GPUMOUNTER_PROBE_MIN_P2P_GBS")) min_p2p_gbs = (double)parse_u32(v);
    std::vector<gmp_device> expected, gone, also;
    std::string dev_dir;
    // environment defaults (the worker's DaemonSet can set them once; command-line flags win)
    if (const char* v = getenv("GPUMOUNTER_PROBE_BYTES")) o.hbm.bytes = parse_size(v);
    if (const char* v = getenv("GPUMOUNTER_PROBE_ITERS")) o.hbm.iters = parse_u32(v);
    if (const char* v = getenv("GPUMOUNTER_PROBE_PATTERNS"))
        o.hbm.pattern_mask = strcmp(v, "all") == 0 ? (1u << GMP_PAT_COUNT) - 1 : parse_u32(v);
    if (const char* v = getenv("GPUMOUNTER_PROBE_SEED")) o.hbm.seed = parse_u64(v, nullptr);
    if (const char* v = getenv("GPUMOUNTER_PROBE_DMA_BYTES")) o.hbm.dma_bytes = parse_size(v);
    if (const char* v = getenv("GPUMOUNTER_PROBE_P2P_BYTES")) o.p2p.bytes = parse_size(v);
    if (const char* v = getenv("GPUMOUNTER_PROBE_SKIP_P2P")) {
        if (parse_u32(v) != 0) o.flags |= GMP_PROBE_F_SKIP_P2P;
    }
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto need = [&](const char* what) -> const char* {
            if (i + 1 >= argc) {
                fprintf(stderr, "%s needs a value\n", what);
                exit(64);
            }
            return argv[++i];
        };
        if (a == "--uuid") add_device(&expected, need("--uuid"));
        else if (a == "--gone") add_device(&gone, need("--gone"));
        else if (a == "--also-mounted") add_device(&also, need("--also-mounted"));
        else if (a == "--uuids") {
            std::string all = need("--uuids");
            size_t b = 0;
            while (b <= all.size()) {
                const size_t c = all.find(',', b);
                const std::string one = all.substr(b, c == std::string::npos ? std::string::npos : c - b);
                if (!one.empty()) add_device(&expected, one);
                if (c == std::string::npos) break;
                b = c + 1;
            }
        } else if (a == "--bytes") o.hbm.bytes = parse_size(need("--bytes"));
        else if (a == "--iters") o.hbm.iters = parse_u32(need("--iters"));
        else if (a == "--patterns") {
            const std::string v = need("--patterns");
            o.hbm.pattern_mask = v == "all" ? (1u << GMP_PAT_COUNT) - 1 : parse_u32(v.c_str());
        } else if (a == "--seed") o.hbm.seed = parse_u64(need("--seed"), nullptr);
        else if (a == "--dma-bytes") o.hbm.dma_bytes = parse_size(need("--dma-bytes"));
        else if (a == "--p2p-bytes") o.p2p.bytes = parse_size(need("--p2p-bytes"));
        else if (a == "--p2p-iters") o.p2p.iters = parse_u32(need("--p2p-iters"));
        else if (a == "--dev-dir") dev_dir = need("--dev-dir");
        else if (a == "--skip-nodes") o.flags |= GMP_PROBE_F_SKIP_NODES;
        else if (a == "--skip-hbm") o.flags |= GMP_PROBE_F_SKIP_HBM;
        else if (a == "--skip-p2p") o.flags |= GMP_PROBE_F_SKIP_P2P;
        else if (a == "--skip-nvml") o.flags |= GMP_PROBE_F_SKIP_NVML;
        else if (a == "--min-hbm-gbs") min_hbm_gbs = (double)parse_u32(need("--min-hbm-gbs"));
        else if (a == "--min-p2p-gbs") min_p2p_gbs = (double)parse_u32(need("--min-p2p-gbs"));
        else if (a == "--strict-health") strict_health = true;
        else if (a == "--nodes-only") o.flags |= GMP_PROBE_F_NODES_ONLY;
        else if (a == "--host-view") o.flags |= GMP_PROBE_F_HOST_VIEW;
        else if (a == "--skip-copy") o.hbm.flags |= GMP_HBM_F_SKIP_COPY;
        else if (a == "--lib-baselines") o.hbm.flags |= GMP_HBM_F_LIB_BASELINES;
        else if (a == "--matrix-only") o.p2p.flags |= GMP_P2P_F_MATRIX_ONLY;
        else if (a == "--tune") {  // benchmarking: --tune knob=value (gmp_tune)
            const std::string kv = need("--tune");
            const size_t eq = kv.find('=');
            if (eq == std::string::npos || gmp_tune(kv.substr(0, eq).c_str(), (int)parse_u32(kv.c_str() + eq + 1)) < 0) {
                fprintf(stderr, "bad --tune '%s': %s\n", kv.c_str(), gmp_last_error());
                return 64;
            }
        }
        else if (a == "-h" || a == "--help") {
            usage();
            return 0;
        } else {
            fprintf(stderr, "unknown argument '%s'\n", a.c_str());
            usage();
            return 64;
        }
    }
    if (!dev_dir.empty()) o.dev_dir = dev_dir.c_str();
    if (!gone.empty()) {
        // RemoveGPU-side mode: only check that the unmounted devices' nodes are gone from this context
        const int rc = gmp_verify_gone(gone.data(), (int)gone.size(), o.dev_dir);
        std::string detail = gmp_last_error();
        for (char& c : detail)
            if (c == '"' || c == '\\') c = '\'';
        printf("{\"status\":%d,\"verdict\":\"%s\",\"detail\":\"%s\",\"checked\":%zu}\n", rc, gmp_strerror(rc), detail.c_str(),
               gone.size());
        fflush(stdout);
        return rc == GMP_OK ? 0 : -rc;
    }
    if (!also.empty()) {
        o.also_visible = also.data();
        o.n_also_visible = (int)also.size();
    }
    if ((int)expected.size() > GMP_MAX_DEVICES || (int)also.size() > GMP_MAX_DEVICES) {
        fprintf(stderr, "at most %d devices\n", GMP_MAX_DEVICES);
        return 64;
    }
    static gmp_report report;  // ~100 KB: keep it off the stack
    int rc = gmp_probe_all(expected.empty() ? nullptr : expected.data(), (int)expected.size(), &o, &report);
    if (rc == GMP_OK) rc = gmp_report_apply_thresholds(&report, min_hbm_gbs, min_p2p_gbs);
    if (rc == GMP_OK && strict_health) rc = gmp_report_check_health(&report);
    static char json[1 << 16];
    if (gmp_report_json(&report, json, sizeof(json)) == GMP_OK) puts(json);
    else printf("{\"status\":%d,\"detail\":\"report too large\"}\n", rc);
    fflush(stdout);
    return rc == GMP_OK ? 0 : -rc;
}
