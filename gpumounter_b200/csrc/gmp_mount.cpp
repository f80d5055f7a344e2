// Native node-local mount path: section 6 of include/gpumounter_probe.h.
//
// Behavioural contract = the reference's Go packages (cited per function); mechanism =
// direct system calls.  The reference spends two to three fork/exec chains per GPU
// (sh -> echo; nsenter -> sh -> mknod) and needs sh/mknod/rm/kill inside the target image
// (ref: docs/guide/FAQ.md:3-4).  Here: open/write for the device-cgroup line and one helper
// process per call (clone(CLONE_VM|CLONE_VFORK) on a private stack) that setns()es into the
// target's mount namespace and issues mknod(2) / unlink(2) itself.  A helper is still needed
// because setns(CLONE_NEWNS) is refused for a multithreaded caller (the worker is one), but it
// never exec()s anything and never copies the worker's page tables.

#include <errno.h>
#include <fcntl.h>
#include <poll.h>
#include <pthread.h>
#include <sched.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <grp.h>
#include <linux/capability.h>
#include <sys/prctl.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <sys/sysmacros.h>
#include <sys/vfs.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <new>
#include <string>
#include <vector>

#include "gmp_nvml_hold.h"
#include "gpumounter_probe.h"

namespace gmp {
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
void clear_error();
}  // namespace gmp

namespace {

using gmp::clear_error;
using gmp::set_error;

constexpr const char* kDevicesRoot = "/sys/fs/cgroup/devices";  // ref: cgroup.go:116
constexpr const char* kDevPrefix = "/dev/nvidia";               // ref: nvidia.go:40
constexpr const char* kCgroupPermission = "rw";                 // ref: nvidia.go:38
constexpr mode_t kDeviceFileMode = 0666;                        // ref: nvidia.go:39 "666"
constexpr const char* kFreeState = "GPU_FREE_STATE";            // ref: nvidia.go:21
constexpr const char* kAllocatedState = "GPU_ALLOCATED_STATE";  // ref: nvidia.go:22
constexpr const char* kGpuPoolNamespace = "gpu-pool";           // ref: types.go:18
constexpr const char* kNvidiaResourceName = "nvidia.com/gpu";   // ref: types.go:10

// Copies into a caller buffer; GMP_ERR_BUFFER_TOO_SMALL when it does not fit.
int emit(const std::string& s, char* out, size_t cap) {
    if (out == nullptr || cap < s.size() + 1) {
        set_error("output buffer of %zu bytes too small for %zu", cap, s.size() + 1);
        return GMP_ERR_BUFFER_TOO_SMALL;
    }
    memcpy(out, s.c_str(), s.size() + 1);
    return GMP_OK;
}

template <size_t N>
void put(char (&dst)[N], const std::string& s) {
    const size_t n = std::min(s.size(), N - 1);
    memcpy(dst, s.data(), n);
    dst[n] = '\0';
}

const char* or_default(const char* v, const char* d) { return (v != nullptr && v[0] != '\0') ? v : d; }

bool extension_on(const gmp_mount_env* env, uint32_t bit) { return env != nullptr && (env->extensions & bit) != 0; }

// Where the pod's cgroup directory lives: the v1 `devices` hierarchy (ref: cgroup.go:116), or, with GMP_EXT_CGROUP_V2
// (section 7 of the header, no reference counterpart), the unified hierarchy.
std::string devices_root(const gmp_mount_env* env) {
    if (extension_on(env, GMP_EXT_CGROUP_V2)) return or_default(env->cgroup2_root, "/sys/fs/cgroup");
    return or_default(env ? env->devices_root : nullptr, kDevicesRoot);
}
std::string dev_prefix(const gmp_mount_env* env) { return or_default(env ? env->dev_prefix : nullptr, kDevPrefix); }
std::string proc_root(const gmp_mount_env* env) { return or_default(env ? env->proc_root : nullptr, "/proc"); }

// The node path of a GPU: the record's DeviceFilePath when the default prefix is in force
// (that is what the reference passes to mknod/rm, ref: namespace.go:168,180), else prefix+minor.
std::string device_node_path(const gmp_mount_env* env, const gmp_gpu* gpu) {
    if (env == nullptr || env->dev_prefix == nullptr || env->dev_prefix[0] == '\0') return gpu->device_file_path;
    return dev_prefix(env) + std::to_string(gpu->minor_number);
}

// ---- resource.Quantity, value only (k8s.io/apimachinery v0.18.6, go.mod:12) -----------------
// Exact value in nano units held in 128 bits: enough for Ei-scale memory and nano-CPU.
using i128 = __int128;

bool parse_quantity(const char* text, i128* nano) {
    if (text == nullptr) return false;
    std::string s(text);
    size_t i = 0;
    bool neg = false;
    if (i < s.size() && (s[i] == '+' || s[i] == '-')) neg = s[i++] == '-';
    i128 mant = 0;
    int ndigits = 0, nfrac = 0;
    for (; i < s.size() && isdigit((unsigned char)s[i]); ++i, ++ndigits) mant = mant * 10 + (s[i] - '0');
    if (i < s.size() && s[i] == '.') {
        for (++i; i < s.size() && isdigit((unsigned char)s[i]); ++i, ++ndigits, ++nfrac) mant = mant * 10 + (s[i] - '0');
    }
    if (ndigits == 0) return false;
    const std::string suffix = s.substr(i);
    int pow10 = -nfrac, pow2 = 0;
    static const struct { const char* sfx; int p10; int p2; } table[] = {
        {"", 0, 0},   {"n", -9, 0}, {"u", -6, 0}, {"m", -3, 0}, {"k", 3, 0},   {"M", 6, 0},   {"G", 9, 0},
        {"T", 12, 0}, {"P", 15, 0}, {"E", 18, 0}, {"Ki", 0, 10}, {"Mi", 0, 20}, {"Gi", 0, 30}, {"Ti", 0, 40},
        {"Pi", 0, 50}, {"Ei", 0, 60}};
    bool known = false;
    for (const auto& t : table) {
        if (suffix == t.sfx) {
            pow10 += t.p10;
            pow2 = t.p2;
            known = true;
            break;
        }
    }
    if (!known) {  // decimal exponent form: e3, E-2
        if (suffix.size() < 2 || (suffix[0] != 'e' && suffix[0] != 'E')) return false;
        size_t j = 1;
        if (suffix[j] == '+' || suffix[j] == '-') ++j;
        if (j >= suffix.size()) return false;
        for (size_t k = j; k < suffix.size(); ++k)
            if (!isdigit((unsigned char)suffix[k])) return false;
        long e = 0;
        for (size_t k = j; k < suffix.size() && e <= 1000; ++k) e = e * 10 + (suffix[k] - '0');
        if (suffix[1] == '-') e = -e;
        if (e > 30 || e < -30) return false;
        pow10 += (int)e;
    }
    i128 v = mant << pow2;
    for (pow10 += 9; pow10 > 0; --pow10) v *= 10;
    for (; pow10 < 0; ++pow10) v = (v + 9) / 10;  // anything finer than nano rounds up, as ParseQuantity does
    *nano = neg ? -v : v;
    return true;
}

// ---- QoS (ref: cgroup.go:177-237) -------------------------------------------------------------
struct ResourceTotals {
    bool present[2] = {false, false};  // [cpu, memory]
    i128 value[2] = {0, 0};
    int count() const { return (int)present[0] + (int)present[1]; }
    void add(int k, i128 q) {
        value[k] = present[k] ? value[k] + q : q;
        present[k] = true;
    }
};

int pod_qos(const gmp_pod* pod, int* qos) {
    ResourceTotals requests, limits;
    bool guaranteed = true;
    for (int c = 0; c < pod->n_containers; ++c) {
        const gmp_container_resources& r = pod->containers[c];
        const char* req[2] = {r.cpu_request, r.memory_request};
        const char* lim[2] = {r.cpu_limit, r.memory_limit};
        bool limit_found[2] = {false, false};
        for (int k = 0; k < 2; ++k) {
            i128 q;
            if (req[k] != nullptr) {
                if (!parse_quantity(req[k], &q)) {
                    set_error("unparsable quantity '%s'", req[k]);
                    return GMP_ERR_INVALID;
                }
                if (q > 0) requests.add(k, q);  // ref: cgroup.go:188-196
            }
            if (lim[k] != nullptr) {
                if (!parse_quantity(lim[k], &q)) {
                    set_error("unparsable quantity '%s'", lim[k]);
                    return GMP_ERR_INVALID;
                }
                if (q > 0) {  // ref: cgroup.go:204-213
                    limit_found[k] = true;
                    limits.add(k, q);
                }
            }
        }
        if (!(limit_found[0] && limit_found[1])) guaranteed = false;  // ref: cgroup.go:216-218
    }
    if (requests.count() == 0 && limits.count() == 0) {
        *qos = GMP_QOS_BESTEFFORT;
        return GMP_OK;
    }
    if (guaranteed) {
        for (int k = 0; k < 2; ++k)
            if (requests.present[k] && !(limits.present[k] && limits.value[k] == requests.value[k])) guaranteed = false;
    }
    *qos = (guaranteed && requests.count() == limits.count()) ? GMP_QOS_GUARANTEED : GMP_QOS_BURSTABLE;
    return GMP_OK;
}

// ---- cgroup names ----------------------------------------------------------------------------
bool component_ok(const std::string& c) { return c.find('/') == std::string::npos && c.find('_') == std::string::npos; }

// runc v1.0.0-rc92 systemd.ExpandSlice (go.mod:8; called at cgroup.go:62)
bool expand_slice(const std::string& slice, std::string* out) {
    static const std::string suffix = ".slice";
    if (slice.size() < suffix.size() || slice.compare(slice.size() - suffix.size(), suffix.size(), suffix) != 0) return false;
    if (slice.find('/') != std::string::npos) return false;
    const std::string name = slice.substr(0, slice.size() - suffix.size());
    if (name == "-") {
        *out = "/";
        return true;
    }
    std::string path, prefix;
    size_t begin = 0;
    while (true) {
        const size_t dash = name.find('-', begin);
        const std::string component = name.substr(begin, dash == std::string::npos ? std::string::npos : dash - begin);
        if (component.empty()) return false;
        path += "/" + prefix + component + suffix;
        prefix += component + "-";
        if (dash == std::string::npos) break;
        begin = dash + 1;
    }
    *out = path;
    return true;
}

int to_systemd(const std::vector<std::string>& parts, std::string* out) {
    if (parts.empty() || (parts.size() == 1 && parts[0].empty())) {
        *out = "/";
        return GMP_OK;
    }
    std::string joined;
    for (size_t i = 0; i < parts.size(); ++i) {
        if (i) joined += '-';
        std::string escaped = parts[i];
        std::replace(escaped.begin(), escaped.end(), '-', '_');  // ref: cgroup.go:70-72
        joined += escaped;
    }
    if (!expand_slice(joined + ".slice", out)) {
        // the reference panics here (cgroup.go:63-66)
        set_error("error converting cgroup name to systemd format: invalid slice name: %s.slice", joined.c_str());
        return GMP_ERR_NAME;
    }
    return GMP_OK;
}

// "/" + path.Join(parts...) for slash-free parts: empties dropped, "." and ".." cleaned
std::string to_cgroupfs(const std::vector<std::string>& parts) {
    std::vector<std::string> kept;
    int up = 0;
    bool any = false;
    for (const std::string& p : parts) {
        if (p.empty()) continue;
        any = true;
        if (p == ".") continue;
        if (p == "..") {
            if (!kept.empty()) kept.pop_back();
            else ++up;
            continue;
        }
        kept.push_back(p);
    }
    if (!any) return "/";
    std::string joined;
    for (int i = 0; i < up; ++i) joined += (joined.empty() ? "" : "/") + std::string("..");
    for (const std::string& k : kept) joined += (joined.empty() ? "" : "/") + k;
    if (joined.empty()) joined = ".";
    return "/" + joined;
}

int cgroup_driver(const gmp_mount_env* env, std::string* out) {
    const char* v = (env != nullptr && env->cgroup_driver != nullptr) ? env->cgroup_driver : getenv("CGROUP_DRIVER");
    const std::string value = v ? v : "";
    if (value != "systemd" && value != "cgroupfs") {
        set_error("unsupported cgroup driver: %s", value.c_str());  // ref: cgroup.go:81
        return GMP_ERR_CGROUP_DRIVER;
    }
    *out = value;
    return GMP_OK;
}

int cgroup_name(const std::string& driver, const gmp_pod* pod, const std::string& container_id, std::string* out,
                const char* scope_prefix = "docker-", const char* cgroupfs_prefix = "") {
    int qos = 0;
    int rc = pod_qos(pod, &qos);
    if (rc != GMP_OK) return rc;
    std::vector<std::string> parts = {"kubepods"};  // ref: cgroup.go:87
    if (qos == GMP_QOS_BURSTABLE) parts.push_back("burstable");
    if (qos == GMP_QOS_BESTEFFORT) parts.push_back("besteffort");
    const std::string pod_part = std::string("pod") + (pod->uid ? pod->uid : "");  // ref: cgroup.go:101
    if (!component_ok(pod_part)) {
        set_error("invalid character in component [\"%s\"] of CgroupName", pod_part.c_str());  // ref: cgroup.go:38
        return GMP_ERR_NAME;
    }
    parts.push_back(pod_part);
    if (driver == "systemd") {
        std::string base;
        rc = to_systemd(parts, &base);
        if (rc != GMP_OK) return rc;
        *out = base + "/" + scope_prefix + container_id + ".scope";  // "docker-", ref: cgroup.go:106
        return GMP_OK;
    }
    if (driver == "cgroupfs") {
        *out = to_cgroupfs(parts) + "/" + cgroupfs_prefix + container_id;  // ref: cgroup.go:108
        return GMP_OK;
    }
    set_error("unsupported cgroup driver");  // ref: cgroup.go:112
    return GMP_ERR_CGROUP_DRIVER;
}

std::string strip_docker_scheme(const char* id) {  // strings.Replace(id, "docker://", "", 1), ref: util.go:23
    std::string s = id ? id : "";
    const size_t at = s.find("docker://");
    if (at != std::string::npos) s.erase(at, strlen("docker://"));
    return s;
}

int pod_cgroup_path(const gmp_mount_env* env, const gmp_pod* pod, std::string* out) {
    if (pod == nullptr || pod->container_id == nullptr) {
        set_error("pod has no container status");  // Go: index out of range on ContainerStatuses[0]
        return GMP_ERR_INVALID;
    }
    std::string driver;
    int rc = cgroup_driver(env, &driver);  // ref: util.go:25-29
    if (rc != GMP_OK) return rc;
    if (extension_on(env, GMP_EXT_CRI_RUNTIMES)) {
        // kubelet reports "<runtime>://<id>"; the reference only knows dockershim.  Leaf names as the runtimes create
        // them: containerd "cri-containerd-<id>.scope" / "<id>", CRI-O "crio-<id>.scope" / "crio-<id>".
        const std::string id = pod->container_id;
        const struct {
            const char* scheme;
            const char* scope_prefix;
            const char* cgroupfs_prefix;
        } runtimes[] = {{"containerd://", "cri-containerd-", ""}, {"cri-o://", "crio-", "crio-"}};
        for (const auto& r : runtimes)
            if (id.compare(0, strlen(r.scheme), r.scheme) == 0)
                return cgroup_name(driver, pod, id.substr(strlen(r.scheme)), out, r.scope_prefix, r.cgroupfs_prefix);
    }
    return cgroup_name(driver, pod, strip_docker_scheme(pod->container_id), out);  // ref: util.go:30-34
}

// ---- cgroup files -------------------------------------------------------------------------------
int read_cgroup_pids(const gmp_mount_env* env, const std::string& cgroup_path, std::vector<std::string>* pids) {
    const std::string file = devices_root(env) + cgroup_path + "/cgroup.procs";
    const int fd = open(file.c_str(), O_RDONLY | O_CLOEXEC);
    if (fd < 0) {
        set_error("open %s: %s", file.c_str(), strerror(errno));
        return GMP_ERR_SYS;
    }
    std::string content;
    char buf[4096];
    ssize_t r;
    while ((r = read(fd, buf, sizeof(buf))) > 0) content.append(buf, (size_t)r);
    close(fd);
    size_t begin = 0;
    for (size_t nl; (nl = content.find('\n', begin)) != std::string::npos; begin = nl + 1)
        pids->push_back(content.substr(begin, nl - begin));
    // whatever follows the last '\n' is dropped, as bufio.ReadString+break does (ref: cgroup.go:131-135)
    return GMP_OK;
}

int write_permission(const gmp_mount_env* env, const std::string& cgroup_path, int minor, const char* file_name) {
    if (extension_on(env, GMP_EXT_CGROUP_V2)) {
        // no devices.allow / devices.deny on the unified hierarchy: edit the device program(s) attached to the cgroup
        gmp_dev_rule rule;
        rule.type = GMP_DEV_CHAR;
        rule.major = GMP_NVIDIA_MAJOR;
        rule.minor = minor;
        rule.access = GMP_DEV_ACC_READ | GMP_DEV_ACC_WRITE;  // "rw", ref: nvidia.go:38
        rule.allow = 1;
        const std::string dir = devices_root(env) + cgroup_path;
        return gmp_cgroup2_update_devices(dir.c_str(), &rule, 1, strcmp(file_name, "devices.allow") == 0, nullptr);
    }
    const std::string file = devices_root(env) + cgroup_path + "/" + file_name;
    // what `echo '...' > file` does: O_WRONLY|O_CREAT|O_TRUNC and a trailing newline
    const int fd = open(file.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0666);
    if (fd < 0) {
        set_error("open %s: %s", file.c_str(), strerror(errno));
        return GMP_ERR_SYS;
    }
    char line[64];
    const int len = snprintf(line, sizeof(line), "c %d:%d %s\n", GMP_NVIDIA_MAJOR, minor, kCgroupPermission);
    const ssize_t w = write(fd, line, (size_t)len);
    const int saved = errno;
    close(fd);
    if (w != len) {
        set_error("write '%.*s' to %s: %s", len - 1, line, file.c_str(), w < 0 ? strerror(saved) : "short write");
        return GMP_ERR_SYS;
    }
    return GMP_OK;
}

// All G GPUs of one request.  v1: one file write per GPU, as the reference's loop does (ref: server.go:80-95 ->
// cgroup.go:143-169).  v2: ONE program replacement carrying all G rules (each replacement costs a verifier run).
int write_permissions(const gmp_mount_env* env, const std::string& cgroup_path, const gmp_gpu* gpus, int n, const char* file_name) {
    if (extension_on(env, GMP_EXT_CGROUP_V2) && n > 1) {
        std::vector<gmp_dev_rule> rules((size_t)n);
        for (int i = 0; i < n; ++i) {
            rules[(size_t)i].type = GMP_DEV_CHAR;
            rules[(size_t)i].major = GMP_NVIDIA_MAJOR;
            rules[(size_t)i].minor = gpus[i].minor_number;
            rules[(size_t)i].access = GMP_DEV_ACC_READ | GMP_DEV_ACC_WRITE;
            rules[(size_t)i].allow = 1;
        }
        const std::string dir = devices_root(env) + cgroup_path;
        return gmp_cgroup2_update_devices(dir.c_str(), rules.data(), n, strcmp(file_name, "devices.allow") == 0, nullptr);
    }
    for (int i = 0; i < n; ++i) {
        const int rc = write_permission(env, cgroup_path, gpus[i].minor_number, file_name);
        if (rc != GMP_OK) return rc;
    }
    return GMP_OK;
}

// ---- namespace work in a forked child ---------------------------------------------------------------
enum class NodeOp { kCreate, kRemove, kOpen, kAbsent };

struct NodeJob {
    NodeOp op;
    std::string path;
    int minor;
};

// A helper process that shares the caller's address space and suspends the caller until it exits or execs
// (clone(CLONE_VM | CLONE_VFORK), the mechanism behind posix_spawn).  A plain fork() of a worker that has a CUDA
// context or a large heap spends ~10 ms just copying page tables (measured: native mount 0.7 ms from a small
// process, 9.6 ms from inside a process that had loaded torch) — more than everything else in the mount.
// The child gets its own stack, its own fd table and its own fs_struct (so setns(CLONE_NEWNS) and umask() stay
// private), and may only make system calls on memory prepared beforehand.
constexpr size_t kHelperStackBytes = 256 * 1024;

struct HelperStack {
    char* base;
    HelperStack() : base(static_cast<char*>(malloc(kHelperStackBytes))) {}
    ~HelperStack() { free(base); }
    void* top() const { return base + kHelperStackBytes; }  // stacks grow down; malloc keeps 16-byte alignment
};

pid_t spawn_helper(int (*fn)(void*), void* arg, const HelperStack& stack) {
    if (stack.base == nullptr) {
        errno = ENOMEM;
        return -1;
    }
    // the child must not run the caller's signal handlers on the shared memory
    sigset_t all, old;
    sigfillset(&all);
    pthread_sigmask(SIG_SETMASK, &all, &old);
    const pid_t pid = clone(fn, stack.top(), CLONE_VM | CLONE_VFORK | SIGCHLD, arg);
    const int saved = errno;
    pthread_sigmask(SIG_SETMASK, &old, nullptr);
    errno = saved;
    return pid;
}

struct NsReport {
    int32_t done;
    int32_t err;
    int32_t stage;  // 0 ok, 1 open ns, 2 setns, 3 mknod, 4 chmod, 5 unlink
};

struct NsHelperArgs {
    const char* ns_path;
    const NodeJob* jobs;
    size_t njobs;
    int report_fd;
    const char* join_procs_file;  // when set: the helper first moves ITSELF into this cgroup (kOpen jobs need the pod's verdict)
};

// Writes "<pid>\n" of the calling process to an open cgroup.procs; system calls only (runs in the clone helper).
bool write_own_pid(int fd) {
    char me[24], rev[24];
    int len = 0, n = 0;
    long v = (long)getpid();
    do {
        rev[n++] = (char)('0' + v % 10);
        v /= 10;
    } while (v > 0);
    while (n > 0) me[len++] = rev[--n];
    me[len++] = '\n';
    return write(fd, me, (size_t)len) == len;
}

int ns_helper_main(void* p) {
    // only raw system calls on memory the parent prepared; no allocation, no locks, no stdio
    const NsHelperArgs* a = static_cast<const NsHelperArgs*>(p);
    NsReport rep = {0, 0, 0};
    const int nsfd = open(a->ns_path, O_RDONLY | O_CLOEXEC);
    int cgfd = -1;
    if (nsfd >= 0 && a->join_procs_file != nullptr) {
        cgfd = open(a->join_procs_file, O_WRONLY | O_APPEND | O_CLOEXEC);
        if (cgfd < 0 || !write_own_pid(cgfd)) {
            rep.err = errno;
            rep.stage = 7;
        }
        if (cgfd >= 0) close(cgfd);
    }
    if (nsfd < 0) {
        rep.err = errno;
        rep.stage = 1;
    } else if (rep.stage != 0) {
        // could not join the pod's cgroup: reported as is
    } else if (setns(nsfd, CLONE_NEWNS) != 0) {
        rep.err = errno;
        rep.stage = 2;
    } else {
        umask(0);
        for (size_t k = 0; k < a->njobs; ++k) {
            const NodeJob& j = a->jobs[k];
            const char* path = j.path.c_str();
            if (j.op == NodeOp::kCreate) {
                if (mknod(path, S_IFCHR | kDeviceFileMode, makedev(GMP_NVIDIA_MAJOR, (unsigned)j.minor)) != 0) {
                    rep.err = errno;
                    rep.stage = 3;
                    break;
                }
                if (chmod(path, kDeviceFileMode) != 0) {  // `mknod -m 666` sets the mode exactly
                    rep.err = errno;
                    rep.stage = 4;
                    break;
                }
            } else if (j.op == NodeOp::kOpen) {
                const int fd = open(path, O_RDWR | O_CLOEXEC);
                if (fd < 0) {
                    rep.err = errno;
                    rep.stage = 6;
                    break;
                }
                close(fd);
            } else if (j.op == NodeOp::kAbsent) {
                struct stat st;
                const int r = lstat(path, &st);
                if (r == 0 || errno != ENOENT) {  // still there (or not even answerable): not "gone"
                    rep.err = r == 0 ? EEXIST : errno;
                    rep.stage = 8;
                    break;
                }
            } else if (unlink(path) != 0) {
                rep.err = errno;
                rep.stage = 5;
                break;
            }
            ++rep.done;
        }
    }
    if (nsfd >= 0) close(nsfd);
    ssize_t ignored = write(a->report_fd, &rep, sizeof(rep));
    (void)ignored;
    _exit(rep.stage == 0 ? 0 : 1);
}

// Runs `jobs` inside the mount namespace named by ns_path.  Returns GMP_OK or GMP_ERR_SYS and, through
// *done, how many jobs completed.  The helper reports (jobs done, errno, failing syscall) over a pipe.
int run_in_mount_ns(const std::string& ns_path, const std::vector<NodeJob>& jobs, int* done, const char* join_procs_file = nullptr,
                    int* failed_errno = nullptr, int* failed_stage = nullptr) {
    using Report = NsReport;
    if (done) *done = 0;
    int pfd[2];
    if (pipe2(pfd, O_CLOEXEC) != 0) {
        set_error("pipe2: %s", strerror(errno));
        return GMP_ERR_SYS;
    }
    HelperStack stack;
    NsHelperArgs args = {ns_path.c_str(), jobs.data(), jobs.size(), pfd[1], join_procs_file};
    const pid_t child = spawn_helper(ns_helper_main, &args, stack);
    if (child < 0) {
        set_error("clone: %s", strerror(errno));
        close(pfd[0]);
        close(pfd[1]);
        return GMP_ERR_SYS;
    }
    close(pfd[1]);
    Report rep = {0, EIO, -1};
    ssize_t got;
    do {
        got = read(pfd[0], &rep, sizeof(rep));
    } while (got < 0 && errno == EINTR);
    close(pfd[0]);
    int status = 0;
    while (waitpid(child, &status, 0) < 0 && errno == EINTR) {
    }
    if (got != (ssize_t)sizeof(rep)) {
        set_error("namespace helper died without a report (wait status 0x%x)", status);
        return GMP_ERR_SYS;
    }
    if (done) *done = rep.done;
    if (failed_errno) *failed_errno = rep.err;
    if (failed_stage) *failed_stage = rep.stage;
    if (rep.stage == 0) return GMP_OK;
    static const char* const what[] = {"", "open", "setns", "mknod", "chmod", "unlink", "open", "join", "still present:"};
    const char* target = rep.stage <= 2 ? ns_path.c_str()
                         : rep.stage == 7 ? (join_procs_file ? join_procs_file : "?")
                                          : (rep.done < (int)jobs.size() ? jobs[(size_t)rep.done].path.c_str() : "?");
    set_error("%s %s: %s", what[rep.stage], target, strerror(rep.err));
    return GMP_ERR_SYS;
}

int mount_ns_path(const gmp_mount_env* env, const gmp_ns_config* cfg, std::string* out) {
    if (cfg == nullptr || cfg->target == 0) {
        set_error("Target must be specified ");  // ref: namespace.go:72
        return GMP_ERR_TARGET;
    }
    if (cfg->mount_file != nullptr && cfg->mount_file[0] != '\0') *out = cfg->mount_file;  // --mount=<file>
    else *out = proc_root(env) + "/" + std::to_string(cfg->target) + "/ns/mnt";
    return GMP_OK;
}

bool go_atoi(const std::string& s, int* out) {  // strconv.Atoi: sign + digits, nothing else
    size_t i = (!s.empty() && (s[0] == '+' || s[0] == '-')) ? 1 : 0;
    if (i >= s.size()) return false;
    for (size_t k = i; k < s.size(); ++k)
        if (!isdigit((unsigned char)s[k])) return false;
    long long v = 0;
    for (size_t k = i; k < s.size(); ++k) {
        v = v * 10 + (s[k] - '0');
        if (v > (long long)INT32_MAX + 1) return false;  // strconv.Atoi: value out of range
    }
    if (s[0] == '-') v = -v;
    if (v > INT32_MAX || v < INT32_MIN) return false;
    *out = (int)v;
    return true;
}

int first_pid(const std::vector<std::string>& pids, int* pid) {
    if (pids.empty()) {
        set_error("cgroup.procs lists no process (the reference indexes pids[0] and panics, util.go:50)");
        return GMP_ERR_EMPTY_CGROUP;
    }
    if (!go_atoi(pids[0], pid)) {
        set_error("Invalid PID: %s", pids[0].c_str());  // ref: util.go:52
        return GMP_ERR_INVALID;
    }
    return GMP_OK;
}

// cgroup.procs ∩ NVML pids for one GPU (ref: util.go:169-195), NVML order kept.
int busy_pids(const std::vector<std::string>& cgroup_pids, const gmp_gpu* gpu, std::vector<std::string>* out) {
    std::vector<gmp_process_info> procs(2048);
    int n = 0;
    int rc = gmp_gpu_running_processes(gpu->uuid, procs.data(), (int)procs.size(), &n);
    if (rc != GMP_OK) return rc;
    for (int i = 0; i < n; ++i) {
        const std::string pid = std::to_string((int)procs[(size_t)i].pid);
        if (std::find(cgroup_pids.begin(), cgroup_pids.end(), pid) != cgroup_pids.end()) out->push_back(pid);
    }
    return GMP_OK;
}

int kill_pids(const std::vector<std::string>& pids) {
    // the reference runs `kill p1 p2 ...` (SIGTERM) with host pids: only the mount namespace was entered
    int failures = 0, last_errno = 0;
    for (const std::string& p : pids) {
        int pid = 0;
        if (!go_atoi(p, &pid) || kill(pid, SIGTERM) != 0) {
            ++failures;
            last_errno = errno;
        }
    }
    if (failures) {
        set_error("kill: %d of %zu processes could not be signalled (%s)", failures, pids.size(), strerror(last_errno));
        return GMP_ERR_SYS;
    }
    return GMP_OK;
}

int mount_many(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n, int* done) {
    if (done) *done = 0;
    std::string cgroup_path;
    int rc = pod_cgroup_path(env, pod, &cgroup_path);
    if (rc != GMP_OK) return rc;
    rc = write_permissions(env, cgroup_path, gpus, n, "devices.allow");  // ref: util.go:37-40
    if (rc != GMP_OK) return rc;
    std::vector<std::string> pids;  // ref: util.go:44-55
    rc = read_cgroup_pids(env, cgroup_path, &pids);
    if (rc != GMP_OK) return rc;
    int pid = 0;
    rc = first_pid(pids, &pid);
    if (rc != GMP_OK) return rc;
    gmp_ns_config cfg;  // ref: util.go:60-63
    memset(&cfg, 0, sizeof(cfg));
    cfg.mount = 1;
    cfg.target = pid;
    std::string ns_path;
    rc = mount_ns_path(env, &cfg, &ns_path);
    if (rc != GMP_OK) return rc;
    std::vector<NodeJob> jobs;
    for (int i = 0; i < n; ++i) jobs.push_back({NodeOp::kCreate, device_node_path(env, &gpus[i]), gpus[i].minor_number});
    return run_in_mount_ns(ns_path, jobs, done);  // ref: util.go:64-67
}

int unmount_many(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n, int force, int* done) {
    if (done) *done = 0;
    std::string cgroup_path;
    int rc = pod_cgroup_path(env, pod, &cgroup_path);
    if (rc != GMP_OK) return rc;
    std::vector<std::string> pids;  // ref: util.go:93-98
    rc = read_cgroup_pids(env, cgroup_path, &pids);
    if (rc != GMP_OK) return rc;
    std::vector<std::string> victims;  // ref: util.go:100-109
    gmp::NvmlHold one_attach_for_all_gpus;
    for (int i = 0; i < n; ++i) {
        std::vector<std::string> busy;
        rc = busy_pids(pids, &gpus[i], &busy);
        if (rc != GMP_OK) return rc;
        if (!busy.empty() && !force) {
            set_error("GPU: %s is busy", gpus[i].device_file_path);
            return GMP_ERR_BUSY;
        }
        for (const std::string& b : busy)
            if (std::find(victims.begin(), victims.end(), b) == victims.end()) victims.push_back(b);
    }
    rc = write_permissions(env, cgroup_path, gpus, n, "devices.deny");  // ref: util.go:112-115
    if (rc != GMP_OK) return rc;
    int pid = 0;  // ref: util.go:118-124
    rc = first_pid(pids, &pid);
    if (rc != GMP_OK) return rc;
    gmp_ns_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.mount = 1;
    cfg.target = pid;
    std::string ns_path;
    rc = mount_ns_path(env, &cfg, &ns_path);
    if (rc != GMP_OK) return rc;
    std::vector<NodeJob> jobs;
    for (int i = 0; i < n; ++i) jobs.push_back({NodeOp::kRemove, device_node_path(env, &gpus[i]), gpus[i].minor_number});
    rc = run_in_mount_ns(ns_path, jobs, done);  // ref: util.go:131-134
    if (rc != GMP_OK) return rc;
    if (!victims.empty()) return kill_pids(victims);  // ref: util.go:137-142
    return GMP_OK;
}

std::vector<std::string> to_vector(const char* const* parts, int n) {
    std::vector<std::string> v;
    for (int i = 0; i < n; ++i) v.emplace_back(parts[i] ? parts[i] : "");
    return v;
}

// encoding/json string in json.Marshal's default (HTML-escaping) mode
std::string json_quote(const char* text) {
    std::string out = "\"";
    for (const unsigned char* c = reinterpret_cast<const unsigned char*>(text); *c; ++c) {
        char esc[8];
        if (*c == '"') out += "\\\"";
        else if (*c == '\\') out += "\\\\";
        else if (*c == '\n') out += "\\n";
        else if (*c == '\r') out += "\\r";
        else if (*c == '\t') out += "\\t";
        else if (*c < 0x20 || *c == '<' || *c == '>' || *c == '&') {
            snprintf(esc, sizeof(esc), "\\u%04x", *c);
            out += esc;
        } else if (*c < 0x80) {
            out += (char)*c;
        } else {
            // Go decodes a rune here: ill-formed UTF-8 is written as \ufffd, LINE/PARAGRAPH SEPARATOR are escaped
            int need = (*c & 0xE0) == 0xC0 ? 1 : (*c & 0xF0) == 0xE0 ? 2 : (*c & 0xF8) == 0xF0 ? 3 : -1;
            uint32_t rune = need == 1 ? (*c & 0x1Fu) : need == 2 ? (*c & 0x0Fu) : (*c & 0x07u);
            for (int k = 1; need > 0 && k <= need; ++k) {
                if ((c[k] & 0xC0) != 0x80) need = -1;
                else rune = (rune << 6) | (c[k] & 0x3Fu);
            }
            static const uint32_t kMin[4] = {0, 0x80, 0x800, 0x10000};
            if (need < 0 || rune < kMin[need] || rune > 0x10FFFF || (rune >= 0xD800 && rune <= 0xDFFF)) {
                out += "\\ufffd";
            } else if (rune == 0x2028 || rune == 0x2029) {
                out += rune == 0x2028 ? "\\u2028" : "\\u2029";
                c += need;
            } else {
                out.append(reinterpret_cast<const char*>(c), (size_t)need + 1);
                c += need;
            }
        }
    }
    return out + "\"";
}

std::vector<int> pod_gpu_indices(const gmp_gpu* list, int n, const std::string& pod_name, const std::string& ns) {
    std::vector<int> idx;
    const std::string needle = pod_name + "-slave-pod-";
    for (int i = 0; i < n; ++i) {
        const std::string owner = list[i].pod_name, owner_ns = list[i].ns;
        const bool own = owner == pod_name && owner_ns == ns;
        const bool slave = owner.find(needle) != std::string::npos && owner_ns == kGpuPoolNamespace;  // substring, ref: collector.go:158
        if (own || slave) idx.push_back(i);
    }
    return idx;
}

int mount_type_of(const gmp_gpu* list, const std::vector<int>& idx, const std::string& pod_name) {
    if (idx.empty()) return GMP_MOUNT_NONE;
    std::vector<std::string> slaves;
    for (int i : idx) {
        const std::string owner = list[i].pod_name;
        if (owner != pod_name && std::find(slaves.begin(), slaves.end(), owner) == slaves.end()) slaves.push_back(owner);
    }
    return slaves.size() < idx.size() ? GMP_MOUNT_ENTIRE : GMP_MOUNT_SINGLE;  // ref: allocator.go:182-186
}

// Every entry point funnels through this so that no C++ exception crosses the C ABI.
template <typename F>
int guarded(F&& body) {
    clear_error();
    try {
        return body();
    } catch (const std::bad_alloc&) {
        set_error("out of host memory");
        return GMP_ERR_NOMEM;
    } catch (...) {
        set_error("unexpected C++ exception");
        return GMP_ERR_INVALID;
    }
}

#define REQUIRE(cond)                          \
    do {                                       \
        if (!(cond)) {                         \
            set_error("bad argument: %s", #cond); \
            return GMP_ERR_INVALID;            \
        }                                      \
    } while (0)

}  // namespace

// ---- C ABI ---------------------------------------------------------------------------------------------
extern "C" int gmp_device_new(int minor, const char* uuid, gmp_gpu* out) {
    return guarded([&]() -> int {
        REQUIRE(out != nullptr && uuid != nullptr);
        memset(out, 0, sizeof(*out));
        out->minor_number = minor;
        put(out->device_file_path, std::string(kDevPrefix) + std::to_string(minor));
        put(out->uuid, uuid);
        put(out->state, kFreeState);
        return GMP_OK;
    });
}

extern "C" int gmp_device_string(const gmp_gpu* g, char* buf, size_t cap) {
    return guarded([&]() -> int {
        REQUIRE(g != nullptr);
        std::string s = "{\"MinorNumber\":" + std::to_string(g->minor_number);
        s += ",\"DeviceFilePath\":" + json_quote(g->device_file_path);
        s += ",\"UUID\":" + json_quote(g->uuid);
        s += ",\"State\":" + json_quote(g->state);
        s += ",\"PodName\":" + json_quote(g->pod_name);
        s += ",\"Namespace\":" + json_quote(g->ns) + "}";
        return emit(s, buf, cap);
    });
}

extern "C" int gmp_device_reset_state(gmp_gpu* g) {
    return guarded([&]() -> int {
        REQUIRE(g != nullptr);
        g->pod_name[0] = '\0';
        g->ns[0] = '\0';
        put(g->state, kFreeState);
        return GMP_OK;
    });
}

extern "C" int gmp_get_pod_qos(const gmp_pod* pod, int* qos) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && qos != nullptr && (pod->n_containers == 0 || pod->containers != nullptr));
        return pod_qos(pod, qos);
    });
}

extern "C" int gmp_cgroup_name_check(const char* component) {
    return guarded([&]() -> int {
        REQUIRE(component != nullptr);
        if (!component_ok(component)) {
            set_error("invalid character in component [\"%s\"] of CgroupName", component);
            return GMP_ERR_NAME;
        }
        return GMP_OK;
    });
}

extern "C" int gmp_cgroup_name_to_systemd(const char* const* parts, int n, char* out, size_t cap) {
    return guarded([&]() -> int {
        REQUIRE(n >= 0 && (n == 0 || parts != nullptr));
        std::string s;
        int rc = to_systemd(to_vector(parts, n), &s);
        return rc != GMP_OK ? rc : emit(s, out, cap);
    });
}

extern "C" int gmp_cgroup_name_to_cgroupfs(const char* const* parts, int n, char* out, size_t cap) {
    return guarded([&]() -> int {
        REQUIRE(n >= 0 && (n == 0 || parts != nullptr));
        return emit(to_cgroupfs(to_vector(parts, n)), out, cap);
    });
}

extern "C" int gmp_get_cgroup_driver(char* out, size_t cap) {
    return guarded([&]() -> int {
        std::string d;
        int rc = cgroup_driver(nullptr, &d);
        if (rc != GMP_OK) {
            if (out && cap) out[0] = '\0';
            return rc;
        }
        return emit(d, out, cap);
    });
}

extern "C" int gmp_get_cgroup_name(const char* driver, const gmp_pod* pod, const char* container_id, char* out, size_t cap) {
    return guarded([&]() -> int {
        REQUIRE(driver != nullptr && pod != nullptr && container_id != nullptr);
        std::string s;
        int rc = cgroup_name(driver, pod, container_id, &s);
        return rc != GMP_OK ? rc : emit(s, out, cap);
    });
}

extern "C" int gmp_detect_cgroup_version(const gmp_mount_env* env, int* version) {
    return guarded([&]() -> int {
        REQUIRE(version != nullptr);
        *version = 0;
        const std::string v1 = or_default(env ? env->devices_root : nullptr, kDevicesRoot);
        struct stat st;
        if (stat(v1.c_str(), &st) == 0 && S_ISDIR(st.st_mode)) {
            *version = 1;
            return GMP_OK;
        }
        const std::string v2 = or_default(env ? env->cgroup2_root : nullptr, "/sys/fs/cgroup");
        struct statfs fs;
        if (statfs(v2.c_str(), &fs) == 0 && (unsigned long)fs.f_type == 0x63677270ul /* CGROUP2_SUPER_MAGIC */) {
            *version = 2;
            return GMP_OK;
        }
        set_error("neither %s (cgroup v1 devices controller) nor a cgroup2 file system at %s", v1.c_str(), v2.c_str());
        return GMP_ERR_UNAVAILABLE;
    });
}

extern "C" int gmp_get_pod_cgroup_path(const gmp_mount_env* env, const gmp_pod* pod, char* out, size_t cap) {
    return guarded([&]() -> int {
        std::string s;
        int rc = pod_cgroup_path(env, pod, &s);
        return rc != GMP_OK ? rc : emit(s, out, cap);
    });
}

extern "C" int gmp_get_device_group_path(const gmp_mount_env* env, const char* cgroup_path, char* out, size_t cap) {
    return guarded([&]() -> int {
        REQUIRE(cgroup_path != nullptr);
        return emit(devices_root(env) + cgroup_path, out, cap);
    });
}

extern "C" int gmp_get_cgroup_pids(const gmp_mount_env* env, const char* cgroup_path, char (*pids)[16], int cap, int* n) {
    return guarded([&]() -> int {
        REQUIRE(cgroup_path != nullptr && n != nullptr && cap >= 0 && (cap == 0 || pids != nullptr));
        *n = 0;
        std::vector<std::string> v;
        int rc = read_cgroup_pids(env, cgroup_path, &v);
        if (rc != GMP_OK) return rc;
        *n = (int)v.size();
        for (int i = 0; i < (int)v.size() && i < cap; ++i) put(pids[i], v[(size_t)i]);
        if ((int)v.size() > cap) {
            set_error("%zu pids, capacity %d", v.size(), cap);
            return GMP_ERR_BUFFER_TOO_SMALL;
        }
        return GMP_OK;
    });
}

extern "C" int gmp_add_gpu_device_permission(const gmp_mount_env* env, const char* cgroup_path, const gmp_gpu* gpu) {
    return guarded([&]() -> int {
        REQUIRE(cgroup_path != nullptr && gpu != nullptr);
        return write_permission(env, cgroup_path, gpu->minor_number, "devices.allow");
    });
}

extern "C" int gmp_remove_gpu_device_permission(const gmp_mount_env* env, const char* cgroup_path, const gmp_gpu* gpu) {
    return guarded([&]() -> int {
        REQUIRE(cgroup_path != nullptr && gpu != nullptr);
        return write_permission(env, cgroup_path, gpu->minor_number, "devices.deny");
    });
}

namespace {
// ref: namespace.go:70-165, flag for flag and in the same order
int build_nsenter_argv(const gmp_ns_config* c, std::vector<std::string>* argv) {
    if (c->target == 0) {
        set_error("Target must be specified ");
        return GMP_ERR_TARGET;
    }
    auto flag = [&](int enabled, const char* file, const char* with_file, const char* bare) {
        if (!enabled) return;
        if (file != nullptr && file[0] != '\0') argv->push_back(std::string(with_file) + file);
        else argv->push_back(bare);
    };
    argv->assign({"nsenter", "--target", std::to_string(c->target)});
    flag(c->cgroup, c->cgroup_file, "--cgroup=", "--cgroup");
    if (c->follow_context) argv->push_back("--follow-context");
    if (c->gid != 0) {
        argv->push_back("--setgid");
        argv->push_back(std::to_string(c->gid));
    }
    flag(c->ipc, c->ipc_file, "--ip=", "--ipc");  // "--ip=" as spelled at namespace.go:96
    flag(c->mount, c->mount_file, "--mount=", "--mount");
    flag(c->net, c->net_file, "--net=", "--net");
    if (c->no_fork) argv->push_back("--no-fork");
    flag(c->pid, c->pid_file, "--pid=", "--pid");
    if (c->preserve_credentials) argv->push_back("--preserve-credentials");
    if (c->root_directory != nullptr && c->root_directory[0] != '\0') {
        argv->push_back("--root");
        argv->push_back(c->root_directory);
    }
    if (c->uid != 0) {
        argv->push_back("--setuid");
        argv->push_back(std::to_string(c->uid));
    }
    flag(c->user, c->user_file, "--user=", "--user");
    flag(c->uts, c->uts_file, "--uts=", "--uts");
    if (c->working_directory != nullptr && c->working_directory[0] != '\0') {
        argv->push_back("--wd");
        argv->push_back(c->working_directory);
    }
    return GMP_OK;
}
}  // namespace

extern "C" int gmp_ns_build_command(const gmp_ns_config* cfg, char* out, size_t cap, int* argc) {
    return guarded([&]() -> int {
        REQUIRE(cfg != nullptr && argc != nullptr);
        std::vector<std::string> argv;
        int rc = build_nsenter_argv(cfg, &argv);
        if (rc != GMP_OK) return rc;
        std::string flat;
        for (const std::string& a : argv) {
            flat += a;
            flat += '\0';
        }
        if (out == nullptr || cap < flat.size()) {
            set_error("output buffer of %zu bytes too small for %zu", cap, flat.size());
            return GMP_ERR_BUFFER_TOO_SMALL;
        }
        memcpy(out, flat.data(), flat.size());
        *argc = (int)argv.size();
        return GMP_OK;
    });
}

extern "C" int gmp_ns_execute(const gmp_ns_config* cfg, const char* program, const char* const* args, int nargs, char* out,
                              size_t out_cap, char* err, size_t err_cap) {
    return guarded([&]() -> int {
        REQUIRE(cfg != nullptr && program != nullptr && nargs >= 0 && (nargs == 0 || args != nullptr));
        std::vector<std::string> argv;
        int rc = build_nsenter_argv(cfg, &argv);
        if (rc != GMP_OK) {
            set_error("Error while building command: Target must be specified  ");  // ref: namespace.go:51
            return rc;
        }
        argv.push_back(program);  // ref: namespace.go:59-60
        for (int i = 0; i < nargs; ++i) argv.push_back(args[i]);
        std::vector<char*> cargv;
        for (std::string& a : argv) cargv.push_back(&a[0]);
        cargv.push_back(nullptr);
        int po[2], pe[2];
        if (pipe2(po, O_CLOEXEC) != 0 || pipe2(pe, O_CLOEXEC) != 0) {
            set_error("pipe2: %s", strerror(errno));
            return GMP_ERR_SYS;
        }
        const pid_t child = fork();
        if (child < 0) {
            set_error("fork: %s", strerror(errno));
            return GMP_ERR_SYS;
        }
        if (child == 0) {
            dup2(po[1], 1);
            dup2(pe[1], 2);
            execvp(cargv[0], cargv.data());
            _exit(127);
        }
        close(po[1]);
        close(pe[1]);
        auto drain = [](int fd, char* dst, size_t cap) {
            size_t len = 0;
            char tmp[1024];
            ssize_t r;
            while ((r = read(fd, tmp, sizeof(tmp))) > 0) {
                if (dst != nullptr && len + 1 < cap) {
                    const size_t take = std::min((size_t)r, cap - 1 - len);
                    memcpy(dst + len, tmp, take);
                    len += take;
                }
            }
            if (dst != nullptr && cap) dst[len] = '\0';
            close(fd);
        };
        drain(po[0], out, out_cap);  // small outputs only: the reference's callers run mknod/rm/kill
        drain(pe[0], err, err_cap);
        int status = 0;
        while (waitpid(child, &status, 0) < 0 && errno == EINTR) {
        }
        if (!WIFEXITED(status) || WEXITSTATUS(status) != 0) {
            set_error("Error while executing command: exit status %d ", WIFEXITED(status) ? WEXITSTATUS(status) : -1);
            return GMP_ERR_SYS;
        }
        return GMP_OK;
    });
}

extern "C" int gmp_ns_add_gpu_device_file(const gmp_mount_env* env, const gmp_ns_config* cfg, const gmp_gpu* gpu) {
    return guarded([&]() -> int {
        REQUIRE(gpu != nullptr);
        std::string ns_path;
        int rc = mount_ns_path(env, cfg, &ns_path);
        if (rc != GMP_OK) return rc;
        return run_in_mount_ns(ns_path, {{NodeOp::kCreate, device_node_path(env, gpu), gpu->minor_number}}, nullptr);
    });
}

extern "C" int gmp_ns_remove_gpu_device_file(const gmp_mount_env* env, const gmp_ns_config* cfg, const gmp_gpu* gpu) {
    return guarded([&]() -> int {
        REQUIRE(gpu != nullptr);
        std::string ns_path;
        int rc = mount_ns_path(env, cfg, &ns_path);
        if (rc != GMP_OK) return rc;
        return run_in_mount_ns(ns_path, {{NodeOp::kRemove, device_node_path(env, gpu), gpu->minor_number}}, nullptr);
    });
}

extern "C" int gmp_ns_kill_running_gpu_processes(const gmp_ns_config* cfg, const char* const* pids, int n) {
    return guarded([&]() -> int {
        REQUIRE(cfg != nullptr && n >= 0 && (n == 0 || pids != nullptr));
        if (cfg->target == 0) {
            set_error("Target must be specified ");
            return GMP_ERR_TARGET;
        }
        return kill_pids(to_vector(pids, n));
    });
}

extern "C" int gmp_can_mount(int mount_type, int request_is_entire_mount) {
    // ref: util.go:207-226
    if (mount_type == GMP_MOUNT_UNKNOWN) return 0;
    if (mount_type != GMP_MOUNT_NONE && request_is_entire_mount) return 0;
    if (mount_type == GMP_MOUNT_ENTIRE) return 0;
    return (mount_type == GMP_MOUNT_NONE || mount_type == GMP_MOUNT_SINGLE) ? 1 : 0;
}

extern "C" int gmp_get_pod_gpu_processes(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpu, char (*pids)[16],
                                         int cap, int* n) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && gpu != nullptr && n != nullptr && cap >= 0 && (cap == 0 || pids != nullptr));
        *n = 0;
        std::string cgroup_path;
        int rc = pod_cgroup_path(env, pod, &cgroup_path);
        if (rc != GMP_OK) return rc;
        std::vector<std::string> in_pod, busy;
        rc = read_cgroup_pids(env, cgroup_path, &in_pod);
        if (rc != GMP_OK) return rc;
        rc = busy_pids(in_pod, gpu, &busy);
        if (rc != GMP_OK) return rc;
        *n = (int)busy.size();
        for (int i = 0; i < (int)busy.size() && i < cap; ++i) put(pids[i], busy[(size_t)i]);
        if ((int)busy.size() > cap) return GMP_ERR_BUFFER_TOO_SMALL;
        return GMP_OK;
    });
}

extern "C" int gmp_mount_gpu(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpu) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && gpu != nullptr);
        return mount_many(env, pod, gpu, 1, nullptr);
    });
}

extern "C" int gmp_unmount_gpu(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpu, int force) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && gpu != nullptr);
        return unmount_many(env, pod, gpu, 1, force, nullptr);
    });
}

extern "C" int gmp_mount_gpus(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n, int* done) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && n >= 0 && (n == 0 || gpus != nullptr));
        if (n == 0) {
            if (done) *done = 0;
            return GMP_OK;
        }
        return mount_many(env, pod, gpus, n, done);
    });
}

extern "C" int gmp_unmount_gpus(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n, int force, int* done) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && n >= 0 && (n == 0 || gpus != nullptr));
        if (n == 0) {
            if (done) *done = 0;
            return GMP_OK;
        }
        return unmount_many(env, pod, gpus, n, force, done);
    });
}

extern "C" int gmp_get_gpu_by_uuid(const gmp_gpu* list, int n, const char* uuid, int* index) {
    return guarded([&]() -> int {
        REQUIRE(uuid != nullptr && index != nullptr && n >= 0 && (n == 0 || list != nullptr));
        for (int i = 0; i < n; ++i) {
            if (strcmp(list[i].uuid, uuid) == 0) {
                *index = i;
                return GMP_OK;
            }
        }
        *index = -1;
        set_error("No GPU with UUID%s", uuid);  // ref: collector.go:87 (no space there)
        return GMP_ERR_NO_DEVICE;
    });
}

extern "C" int gmp_update_gpu_status(gmp_gpu* list, int n, const gmp_pod_resources* pods, int n_pods) {
    return guarded([&]() -> int {
        REQUIRE(n >= 0 && n_pods >= 0 && (n == 0 || list != nullptr) && (n_pods == 0 || pods != nullptr));
        for (int i = 0; i < n; ++i) {  // resetGPUStatus, ref: collector.go:140-144
            list[i].pod_name[0] = '\0';
            list[i].ns[0] = '\0';
            put(list[i].state, kFreeState);
        }
        for (int p = 0; p < n_pods; ++p) {
            for (int c = 0; c < pods[p].n_containers; ++c) {
                const gmp_container_entry& ce = pods[p].containers[c];
                for (int d = 0; d < ce.n_devices; ++d) {
                    const gmp_container_devices& dev = ce.devices[d];
                    if (dev.resource_name == nullptr || strcmp(dev.resource_name, kNvidiaResourceName) != 0) continue;
                    for (int u = 0; u < dev.n_ids; ++u) {
                        int idx = -1;
                        for (int i = 0; i < n && idx < 0; ++i)
                            if (strcmp(list[i].uuid, dev.device_ids[u]) == 0) idx = i;
                        if (idx < 0) {
                            set_error("No GPU with UUID%s", dev.device_ids[u]);
                            return GMP_ERR_NO_DEVICE;  // ref: collector.go:123-125
                        }
                        put(list[idx].state, kAllocatedState);
                        put(list[idx].pod_name, pods[p].name ? pods[p].name : "");
                        put(list[idx].ns, pods[p].ns ? pods[p].ns : "");
                    }
                }
            }
        }
        return GMP_OK;
    });
}

extern "C" int gmp_get_pod_gpu_resources(const gmp_gpu* list, int n, const char* pod_name, const char* ns, int32_t* indices,
                                         int* count) {
    return guarded([&]() -> int {
        REQUIRE(pod_name != nullptr && ns != nullptr && count != nullptr && n >= 0 && (n == 0 || (list != nullptr && indices != nullptr)));
        const std::vector<int> idx = pod_gpu_indices(list, n, pod_name, ns);
        for (size_t i = 0; i < idx.size(); ++i) indices[i] = idx[i];
        *count = (int)idx.size();
        return GMP_OK;
    });
}

extern "C" int gmp_get_mount_type(const gmp_gpu* list, int n, const char* pod_name, const char* ns, int* mount_type) {
    return guarded([&]() -> int {
        REQUIRE(pod_name != nullptr && ns != nullptr && mount_type != nullptr && n >= 0 && (n == 0 || list != nullptr));
        *mount_type = mount_type_of(list, pod_gpu_indices(list, n, pod_name, ns), pod_name);
        return GMP_OK;
    });
}

extern "C" int gmp_get_remove_gpu(const gmp_gpu* list, int n, const char* owner_name, const char* owner_ns,
                                  const char* const* uuids, int n_uuids, int32_t* indices, int* count) {
    return guarded([&]() -> int {
        REQUIRE(owner_name != nullptr && owner_ns != nullptr && count != nullptr && n >= 0 && n_uuids >= 0);
        REQUIRE((n == 0 || (list != nullptr && indices != nullptr)) && (n_uuids == 0 || uuids != nullptr));
        const std::vector<int> mine = pod_gpu_indices(list, n, owner_name, owner_ns);
        const int type = mount_type_of(list, mine, owner_name);
        const std::vector<std::string> wanted = to_vector(uuids, n_uuids);
        std::vector<int> picked;
        for (int i : mine) {  // ref: allocator.go:112-119
            const bool listed = std::find(wanted.begin(), wanted.end(), std::string(list[i].uuid)) != wanted.end();
            if ((type == GMP_MOUNT_ENTIRE || listed) && std::string(list[i].pod_name) != owner_name) picked.push_back(i);
        }
        if ((size_t)n_uuids != picked.size()) picked.clear();  // ref: allocator.go:121-123
        for (size_t i = 0; i < picked.size(); ++i) indices[i] = picked[i];
        *count = (int)picked.size();
        return GMP_OK;
    });
}

extern "C" int gmp_slave_pod_name(const char* owner, const unsigned char rand3[3], char* out, size_t cap) {
    return guarded([&]() -> int {
        REQUIRE(owner != nullptr && rand3 != nullptr);
        char hex[8];
        snprintf(hex, sizeof(hex), "%02x%02x%02x", rand3[0], rand3[1], rand3[2]);
        return emit(std::string(owner) + "-slave-pod-" + hex, out, cap);
    });
}

namespace {
struct ProbeHelperArgs {
    const char* exe;
    const char* ns_path;
    const char* procs_file;
    char* const* argv;
    int stdout_fd;
    // privilege drop (unless GMP_EXT_PROBE_KEEP_PRIVILEGES): the pod's other namespaces and its first process's identity
    int drop;
    const char* extra_ns[3];  // net, ipc, uts of the target; entered when they differ from ours
    unsigned uid, gid;
};

// Child of gmp_probe_in_pod (shares the parent's memory until fexecve): system calls only.
// Stage codes travel back as exit statuses 120..127 so the parent can name the failing step.
int probe_helper_main(void* p) {
    const ProbeHelperArgs* a = static_cast<const ProbeHelperArgs*>(p);
    // not O_CLOEXEC: harmless for the ELF probe (it inherits a descriptor to its own image) and required
    // if the file is a script, whose interpreter is handed /dev/fd/N
    const int exe_fd = open(a->exe, O_RDONLY);
    if (exe_fd < 0) _exit(120);
    const int ns_fd = open(a->ns_path, O_RDONLY | O_CLOEXEC);
    if (ns_fd < 0) _exit(121);
    int extra_fd[3] = {-1, -1, -1};
    if (a->drop)
        for (int k = 0; k < 3; ++k)
            if (a->extra_ns[k] != nullptr) extra_fd[k] = open(a->extra_ns[k], O_RDONLY | O_CLOEXEC);  // absent: stay where we are
    const int cg_fd = open(a->procs_file, O_WRONLY | O_APPEND | O_CLOEXEC);
    if (cg_fd < 0) _exit(122);
    if (!write_own_pid(cg_fd)) _exit(123);
    close(cg_fd);
    for (int k = 0; k < 3; ++k)
        if (extra_fd[k] >= 0 && setns(extra_fd[k], 0) != 0) _exit(126);
    if (setns(ns_fd, CLONE_NEWNS) != 0) _exit(124);
    if (a->drop) {
        // From here on the process may hold nothing the tenant does not: whatever the dynamic loader pulls in from the
        // pod's root (PT_INTERP, libc, libcuda.so.1, libnvidia-ml.so.1) runs with the pod user's identity, no
        // capability to gain any back, and no way to acquire one through a set-uid file.
        for (int cap = 0; cap <= 63; ++cap) (void)prctl(PR_CAPBSET_DROP, cap, 0, 0, 0);  // EINVAL past the kernel's last one
        // RAW system calls: glibc's setuid()/setgid()/setgroups() wrappers broadcast the change to every thread they find
        // in the thread list - which this CLONE_VM child shares with the worker.  The raw calls change this task alone.
        if (syscall(SYS_setgroups, 0, nullptr) != 0) _exit(127);
        if (syscall(SYS_setgid, a->gid) != 0 || syscall(SYS_setuid, a->uid) != 0) _exit(127);
        struct __user_cap_header_struct hdr = {_LINUX_CAPABILITY_VERSION_3, 0};
        struct __user_cap_data_struct none[2] = {{0, 0, 0}, {0, 0, 0}};
        if (syscall(SYS_capset, &hdr, none) != 0) _exit(127);  // uid 0 keeps its sets across setuid(0): clear them by hand
        if (prctl(PR_SET_NO_NEW_PRIVS, 1, 0, 0, 0) != 0) _exit(127);
    }
    dup2(a->stdout_fd, 1);
    sigset_t none_blocked;  // spawn_helper blocked every signal around clone(); the probe must not inherit that
    sigemptyset(&none_blocked);
    sigprocmask(SIG_SETMASK, &none_blocked, nullptr);
    fexecve(exe_fd, a->argv, ::environ);
    _exit(125);
}

// Real uid / gid of a process (the "Uid:" / "Gid:" lines of /proc/<pid>/status, first column).
bool identity_of(const std::string& proc_root_dir, int pid, unsigned* uid, unsigned* gid) {
    const std::string file = proc_root_dir + "/" + std::to_string(pid) + "/status";
    FILE* f = fopen(file.c_str(), "re");
    if (f == nullptr) return false;
    char line[256];
    bool have_uid = false, have_gid = false;
    auto first_number = [](const char* p, unsigned* out) {  // no sscanf/strtoul: glibc redirects them to 2.38-only symbols
        while (*p == ' ' || *p == '\t') ++p;
        if (*p < '0' || *p > '9') return false;
        unsigned v = 0;
        for (; *p >= '0' && *p <= '9'; ++p) v = v * 10u + (unsigned)(*p - '0');
        *out = v;
        return true;
    };
    while (fgets(line, sizeof(line), f) != nullptr) {
        if (strncmp(line, "Uid:", 4) == 0) have_uid = first_number(line + 4, uid);
        else if (strncmp(line, "Gid:", 4) == 0) have_gid = first_number(line + 4, gid);
    }
    fclose(f);
    return have_uid && have_gid;
}

// <proc>/<pid>/ns/<name> when it names another namespace than the caller's own, else "".
std::string foreign_ns(const std::string& proc_root_dir, int pid, const char* name) {
    const std::string theirs = proc_root_dir + "/" + std::to_string(pid) + "/ns/" + name;
    const std::string ours = std::string("/proc/self/ns/") + name;
    char a[64], b[64];
    const ssize_t na = readlink(theirs.c_str(), a, sizeof(a) - 1), nb = readlink(ours.c_str(), b, sizeof(b) - 1);
    if (na <= 0 || nb <= 0) return "";
    a[na] = b[nb] = '\0';
    return strcmp(a, b) == 0 ? "" : theirs;
}
}  // namespace

// ---- the pod's view of the device nodes, without executing anything in the pod ------------------------------
// Stage 1 of the two-stage hook (gmp_probe_all(GMP_PROBE_F_NODES_ONLY) is the same check done by the probe binary): a
// clone helper joins the pod's device cgroup, enters its mount namespace and open(2)s /dev/nvidiactl, /dev/nvidia-uvm
// and every mounted GPU's node.  Nothing of the pod's image is run, so it also works for musl, distroless and scratch
// images, where the probe binary cannot start.
extern "C" int gmp_check_nodes_in_pod(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n, int* failed_index) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && n >= 0 && (n == 0 || gpus != nullptr));
        if (failed_index) *failed_index = -1;
        std::string cgroup_path;
        int rc = pod_cgroup_path(env, pod, &cgroup_path);
        if (rc != GMP_OK) return rc;
        std::vector<std::string> pids;
        rc = read_cgroup_pids(env, cgroup_path, &pids);
        if (rc != GMP_OK) return rc;
        int target = 0;
        rc = first_pid(pids, &target);
        if (rc != GMP_OK) return rc;
        const std::string ns_path = proc_root(env) + "/" + std::to_string(target) + "/ns/mnt";
        const std::string procs_file = devices_root(env) + cgroup_path + "/cgroup.procs";
        // with a custom prefix ("/x/dev/nvidia") the control nodes sit next to the GPU nodes, as in the harness
        const std::string prefix = dev_prefix(env);
        std::vector<NodeJob> jobs = {{NodeOp::kOpen, prefix + "ctl", -1}, {NodeOp::kOpen, prefix + "-uvm", -1}};
        for (int i = 0; i < n; ++i) jobs.push_back({NodeOp::kOpen, device_node_path(env, &gpus[i]), gpus[i].minor_number});
        int done = 0, err = 0, stage = 0;
        rc = run_in_mount_ns(ns_path, jobs, &done, procs_file.c_str(), &err, &stage);
        if (rc == GMP_OK) return GMP_OK;
        if (stage != 6) return rc;  // set-up failure (namespace, cgroup): GMP_ERR_SYS with the step named
        const std::string detail = gmp_last_error();
        set_error("%s", detail.c_str());
        if (done == 0) return GMP_ERR_NODE_CTL;
        if (done == 1) return GMP_ERR_NODE_UVM;
        if (failed_index) *failed_index = done - 2;
        return (err == EPERM || err == EACCES) ? GMP_ERR_NODE_DENIED : GMP_ERR_NODE_GPU;
    });
}

// After an unmount: none of the GPUs' nodes may exist in the pod any more (what `gpumounter-probe --gone` checks from an
// exec'ed binary, here from the helper: nothing of the pod's image runs).
extern "C" int gmp_verify_gone_in_pod(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n, int* failed_index) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && n >= 0 && (n == 0 || gpus != nullptr));
        if (failed_index) *failed_index = -1;
        std::string cgroup_path;
        int rc = pod_cgroup_path(env, pod, &cgroup_path);
        if (rc != GMP_OK) return rc;
        std::vector<std::string> pids;
        rc = read_cgroup_pids(env, cgroup_path, &pids);
        if (rc != GMP_OK) return rc;
        int target = 0;
        rc = first_pid(pids, &target);
        if (rc != GMP_OK) return rc;
        const std::string ns_path = proc_root(env) + "/" + std::to_string(target) + "/ns/mnt";
        std::vector<NodeJob> jobs;
        for (int i = 0; i < n; ++i) jobs.push_back({NodeOp::kAbsent, device_node_path(env, &gpus[i]), gpus[i].minor_number});
        int done = 0, err = 0, stage = 0;
        rc = run_in_mount_ns(ns_path, jobs, &done, nullptr, &err, &stage);
        if (rc == GMP_OK || stage != 8) return rc;
        if (failed_index) *failed_index = done;
        const std::string detail = gmp_last_error();
        set_error("%s", detail.c_str());
        return GMP_ERR_STILL_MOUNTED;
    });
}

// ---- in-pod execution of the probe -------------------------------------------------------------------
// A probe run from the worker's own context proves nothing: the worker is privileged and sees every GPU
// (ref: deploy/gpu-mounter-workers.yaml:23-24).  And `nsenter --mount` alone, which is all the reference
// ever does (ref: namespace.go:102-108), does not move a process into the pod's DEVICES CGROUP, the thing
// that actually gates open("/dev/nvidiaN").  So the child here (1) opens the probe binary while still in the
// worker's namespace, (2) writes itself into the target's device cgroup, (3) setns()es into the target's
// mount namespace and (4) fexecve()s the already open binary: libcuda.so.1 / libnvidia-ml.so.1 then resolve
// from the pod's own root, the device nodes are the pod's, and the cgroup verdict is the pod's.
namespace {

// Where and as whom probe processes for one pod are started: resolved once per call.
struct PodContext {
    std::string cgroup_path, cgroup_dir, ns_path;
    int target = 0;
    int drop = 0;
    unsigned uid = 0, gid = 0;
    std::string extra_ns[3];
};

int resolve_pod_context(const gmp_mount_env* env, const gmp_pod* pod, PodContext* ctx) {
    int rc = pod_cgroup_path(env, pod, &ctx->cgroup_path);
    if (rc != GMP_OK) return rc;
    std::vector<std::string> pids;
    rc = read_cgroup_pids(env, ctx->cgroup_path, &pids);
    if (rc != GMP_OK) return rc;
    rc = first_pid(pids, &ctx->target);
    if (rc != GMP_OK) return rc;
    ctx->ns_path = proc_root(env) + "/" + std::to_string(ctx->target) + "/ns/mnt";
    ctx->cgroup_dir = devices_root(env) + ctx->cgroup_path;
    if (!extension_on(env, GMP_EXT_PROBE_KEEP_PRIVILEGES)) {
        ctx->drop = 1;
        if (!identity_of(proc_root(env), ctx->target, &ctx->uid, &ctx->gid)) {
            set_error("in-pod probe: cannot read the identity of pid %d", ctx->target);
            return GMP_ERR_SYS;
        }
        static const char* const names[3] = {"net", "ipc", "uts"};
        for (int k = 0; k < 3; ++k) ctx->extra_ns[k] = foreign_ns(proc_root(env), ctx->target, names[k]);
    }
    return GMP_OK;
}

struct ProbeChild {
    pid_t pid = -1;
    int fd = -1;  // read end of the child's stdout
    std::string output;
    int wait_status = 0;
};

// Starts one probe process: member of the device cgroup whose cgroup.procs is `procs_file`, inside the pod's namespaces.
int spawn_probe(const PodContext& ctx, const std::string& procs_file, const char* probe_exe, std::vector<std::string>& args,
                ProbeChild* child) {
    std::vector<char*> argv;
    for (std::string& a : args) argv.push_back(&a[0]);
    argv.push_back(nullptr);
    int out_pipe[2];
    if (pipe2(out_pipe, O_CLOEXEC) != 0) {
        set_error("pipe2: %s", strerror(errno));
        return GMP_ERR_SYS;
    }
    HelperStack stack;
    ProbeHelperArgs hargs = {probe_exe, ctx.ns_path.c_str(), procs_file.c_str(), argv.data(), out_pipe[1], ctx.drop,
                             {nullptr, nullptr, nullptr}, ctx.uid, ctx.gid};
    for (int k = 0; k < 3; ++k) hargs.extra_ns[k] = ctx.extra_ns[k].empty() ? nullptr : ctx.extra_ns[k].c_str();
    const pid_t pid = spawn_helper(probe_helper_main, &hargs, stack);  // returns once the child has exec'ed or exited
    const int saved = errno;
    close(out_pipe[1]);
    if (pid < 0) {
        set_error("clone: %s", strerror(saved));
        close(out_pipe[0]);
        return GMP_ERR_SYS;
    }
    child->pid = pid;
    child->fd = out_pipe[0];
    return GMP_OK;
}

// Reads every child's stdout to EOF (a pipe's write end closes when its probe exits), kills all of them when
// timeout_ms (0 = none) runs out, reaps them.  *timed_out says which way it ended.
void collect_probes(std::vector<ProbeChild>& kids, uint32_t timeout_ms, bool* timed_out) {
    *timed_out = false;
    struct timespec t0;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    size_t open_pipes = 0;
    for (const ProbeChild& k : kids)
        if (k.fd >= 0) ++open_pipes;
    char tmp[4096];
    while (open_pipes > 0) {
        int wait_ms = -1;
        if (timeout_ms != 0) {
            struct timespec now;
            clock_gettime(CLOCK_MONOTONIC, &now);
            const long long elapsed = (now.tv_sec - t0.tv_sec) * 1000ll + (now.tv_nsec - t0.tv_nsec) / 1000000ll;
            if (elapsed >= (long long)timeout_ms) {
                *timed_out = true;
                break;
            }
            wait_ms = (int)((long long)timeout_ms - elapsed);
        }
        std::vector<struct pollfd> pf;
        std::vector<size_t> who;
        for (size_t i = 0; i < kids.size(); ++i)
            if (kids[i].fd >= 0) {
                pf.push_back({kids[i].fd, POLLIN, 0});
                who.push_back(i);
            }
        const int pr = poll(pf.data(), pf.size(), wait_ms);
        if (pr < 0 && errno != EINTR) break;
        if (pr <= 0) continue;  // 0: the loop head notices the deadline
        for (size_t k = 0; k < pf.size(); ++k) {
            if (!(pf[k].revents & (POLLIN | POLLHUP | POLLERR))) continue;
            ProbeChild& c = kids[who[k]];
            const ssize_t r = read(c.fd, tmp, sizeof(tmp));
            if (r > 0) {
                c.output.append(tmp, (size_t)r);
            } else if (r == 0 || errno != EINTR) {
                close(c.fd);
                c.fd = -1;
                --open_pipes;
            }
        }
    }
    for (ProbeChild& c : kids) {
        if (*timed_out && c.pid > 0) (void)kill(c.pid, SIGKILL);  // host pid: the pid namespace was not entered
        if (c.fd >= 0) {
            close(c.fd);
            c.fd = -1;
        }
        if (c.pid > 0)
            while (waitpid(c.pid, &c.wait_status, 0) < 0 && errno == EINTR) {
            }
    }
}

// The probe's last output line and its verdict.  GMP_OK / the GMP_ERR_* the probe exited with / GMP_ERR_SYS with the
// failing set-up step named.
int probe_verdict(ProbeChild& c, std::string* line, int* exit_code) {
    const int code = WIFEXITED(c.wait_status) ? WEXITSTATUS(c.wait_status) : -1;
    if (exit_code) *exit_code = code;
    if (code >= 120 && code <= 127) {
        static const char* const stage[] = {"open the probe binary", "open the target's mount namespace",
                                            "open the target's cgroup.procs", "join the target's device cgroup",
                                            "setns into the target's mount namespace", "fexecve the probe",
                                            "enter the target's net / ipc / uts namespace",
                                            "drop to the pod user's identity and an empty capability set"};
        set_error("in-pod probe: could not %s", stage[code - 120]);
        return GMP_ERR_SYS;
    }
    std::string& output = c.output;
    while (!output.empty() && (output.back() == '\n' || output.back() == '\r')) output.pop_back();
    const size_t last_line = output.rfind('\n');
    *line = last_line == std::string::npos ? output : output.substr(last_line + 1);
    if (code < 0) {
        set_error("in-pod probe was killed (wait status 0x%x)", c.wait_status);
        return GMP_ERR_SYS;
    }
    return code == 0 ? GMP_OK : -code;  // the probe's exit status is -(GMP_ERR_*)
}

}  // namespace

extern "C" int gmp_probe_in_pod(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n,
                                const char* probe_exe, const char* const* extra_args, int n_extra, char* json, size_t cap,
                                int* exit_status) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && probe_exe != nullptr && json != nullptr && cap > 0 && n >= 0 && (n == 0 || gpus != nullptr));
        REQUIRE(n_extra >= 0 && (n_extra == 0 || extra_args != nullptr));
        json[0] = '\0';
        if (exit_status) *exit_status = -1;
        PodContext ctx;
        int rc = resolve_pod_context(env, pod, &ctx);
        if (rc != GMP_OK) return rc;
        std::vector<std::string> args = {"gpumounter-probe"};
        for (int i = 0; i < n; ++i) {
            args.push_back("--uuid");
            args.push_back(std::string(gpus[i].uuid) + ":" + std::to_string(gpus[i].minor_number));
        }
        for (int i = 0; i < n_extra; ++i) args.push_back(extra_args[i]);
        std::vector<ProbeChild> kids(1);
        rc = spawn_probe(ctx, ctx.cgroup_dir + "/cgroup.procs", probe_exe, args, &kids[0]);
        if (rc != GMP_OK) return rc;
        const uint32_t timeout_ms = env ? env->probe_timeout_ms : 0;
        bool timed_out = false;
        collect_probes(kids, timeout_ms, &timed_out);
        if (timed_out) {
            set_error("in-pod probe did not finish within %u ms and was killed", timeout_ms);
            return GMP_ERR_TIMEOUT;
        }
        std::string line;
        int code = -1;
        rc = probe_verdict(kids[0], &line, &code);
        if (exit_status) *exit_status = code;
        if (rc == GMP_ERR_SYS && (code >= 120 || code < 0) && line.empty()) return rc;
        const int erc = emit(line, json, cap);
        if (erc != GMP_OK) return erc;
        return rc;
    });
}

// ---- fan-out: one probe process per mounted GPU --------------------------------------------------------------------
// Measured on an 8-GPU HGX B200 (profiles/r02_ctx_timeline_8gpu.json): ONE process that starts CUDA on 8 freshly mounted
// GPUs needs 1.3-1.5 s before every GPU has run a verified kernel - cuInit attaches all eight (~0.55-0.9 s) and the eight
// primary contexts then contend inside the driver (~0.6 s).  Eight processes that each see all eight GPUs are worse (3.9 s).
// Eight processes that each can OPEN only their own GPU take 0.86 s: each cuInit attaches one GPU, and they run side by
// side.  The narrowing is done with the device cgroup itself: a child of the pod's devices group can only hold a subset
// of the pod's permissions, so a kernel that runs and verifies in the child is a fortiori allowed in the pod.
namespace {

bool write_file(const std::string& path, const std::string& text) {
    const int fd = open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0644);  // as write_permission() opens them
    if (fd < 0) return false;
    const ssize_t w = write(fd, text.data(), text.size());
    const int saved = errno;
    close(fd);
    errno = saved;
    return w == (ssize_t)text.size();
}

// "key":<unsigned> at the top level of a report line (both keys used here occur once per report).
unsigned long long json_u64(const std::string& line, const char* key) {
    const std::string needle = std::string("\"") + key + "\":";
    const size_t at = line.rfind(needle);
    if (at == std::string::npos) return 0;
    unsigned long long v = 0;
    for (size_t i = at + needle.size(); i < line.size() && line[i] >= '0' && line[i] <= '9'; ++i) v = v * 10 + (unsigned)(line[i] - '0');
    return v;
}

}  // namespace

extern "C" int gmp_probe_in_pod_fanout(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n,
                                       const char* probe_exe, const char* const* extra_args, int n_extra, char* json,
                                       size_t cap, int* exit_status) {
    return guarded([&]() -> int {
        REQUIRE(pod != nullptr && probe_exe != nullptr && json != nullptr && cap > 0 && n >= 1 && n <= GMP_MAX_DEVICES && gpus != nullptr);
        REQUIRE(n_extra >= 0 && (n_extra == 0 || extra_args != nullptr));
        json[0] = '\0';
        if (exit_status) *exit_status = -1;
        if (extension_on(env, GMP_EXT_CGROUP_V2)) {
            set_error("fan-out narrows each probe with a cgroup-v1 devices child group; not available on the unified hierarchy");
            return GMP_ERR_UNSUPPORTED;
        }
        PodContext ctx;
        int rc = resolve_pod_context(env, pod, &ctx);
        if (rc != GMP_OK) return rc;
        const uint32_t timeout_ms = env ? env->probe_timeout_ms : 0;

        // 1. one child group per GPU, narrowed to this GPU and the control node
        std::vector<std::string> dirs;
        bool not_narrowed = false;  // the pod holds a wildcard over the NVIDIA major: children see what the pod sees
        auto remove_dirs = [&]() {
            for (const std::string& d : dirs)
                for (int attempt = 0; attempt < 200 && rmdir(d.c_str()) != 0; ++attempt) {
                    if (errno == EBUSY) {  // cgroupfs: the probe's exit has not been accounted yet
                        usleep(1000);
                    } else if (errno == ENOTEMPTY && attempt == 0) {  // a plain directory tree (scratch harness): our own files
                        for (const char* f : {"/cgroup.procs", "/devices.deny", "/devices.allow"}) (void)unlink((d + f).c_str());
                    } else {
                        break;
                    }
                }
        };
        static std::atomic<unsigned> call_counter{0};  // concurrent calls for one pod must not share a child group
        const unsigned call_id = call_counter.fetch_add(1, std::memory_order_relaxed);
        for (int i = 0; i < n; ++i) {
            const std::string d = ctx.cgroup_dir + "/gmp-probe-" + std::to_string((long)getpid()) + "-" + std::to_string(call_id) + "-" + std::to_string(i);
            if (mkdir(d.c_str(), 0755) != 0) {
                set_error("mkdir %s: %s", d.c_str(), strerror(errno));
                remove_dirs();
                return GMP_ERR_SYS;
            }
            dirs.push_back(d);
            {   // cgroupfs provides cgroup.procs; a plain directory tree (the harness's scratch shape) needs one made
                const int pfd = open((d + "/cgroup.procs").c_str(), O_WRONLY | O_CREAT | O_CLOEXEC, 0644);
                if (pfd >= 0) close(pfd);
            }
            // A new devices group starts with a copy of its parent's list.  Take every OTHER GPU out of it, one entry at
            // a time: in a default-deny group the kernel revokes exact entries only (a wildcard deny does not remove
            // "c 195:3 rw").  GPUs the pod owns natively are entries like any other and go too.  A pod that holds the
            // wildcard "c 195:* rw" cannot be narrowed this way; its probes then each attach every GPU (slower, still right).
            std::string list;
            {
                const int lfd = open((d + "/devices.list").c_str(), O_RDONLY | O_CLOEXEC);
                char buf[4096];
                ssize_t r;
                while (lfd >= 0 && (r = read(lfd, buf, sizeof(buf))) > 0) list.append(buf, (size_t)r);
                if (lfd >= 0) close(lfd);
            }
            const std::string prefix = "c " + std::to_string(GMP_NVIDIA_MAJOR) + ":";
            bool ok = true;
            size_t begin = 0;
            for (size_t nl; ok && (nl = list.find('\n', begin)) != std::string::npos; begin = nl + 1) {
                const std::string line = list.substr(begin, nl - begin);
                // a wildcard that covers the NVIDIA major with read or write access ("a *:* rwm", "c *:* rw", "c 195:* rw"):
                // this group cannot be narrowed, its probes will see every GPU the pod sees
                const bool any_dev = line.compare(0, 6, "a *:* ") == 0 || line.compare(0, 6, "c *:* ") == 0 ||
                                     line.compare(0, prefix.size() + 2, prefix + "* ") == 0;
                if (any_dev && line.find_first_of("rw", line.rfind(' ') == std::string::npos ? 0 : line.rfind(' ')) != std::string::npos)
                    not_narrowed = true;
                if (line.compare(0, prefix.size(), prefix) != 0) continue;
                int minor = 0;
                size_t k = prefix.size();
                if (k >= line.size() || !isdigit((unsigned char)line[k])) continue;  // "c 195:* ..."
                for (; k < line.size() && isdigit((unsigned char)line[k]); ++k) minor = minor * 10 + (line[k] - '0');
                if (minor == gpus[i].minor_number || minor == 255) continue;
                ok = write_file(d + "/devices.deny", line);
            }
            if (!ok) {
                set_error("narrowing %s to GPU minor %d: %s", d.c_str(), gpus[i].minor_number, strerror(errno));
                remove_dirs();
                return GMP_ERR_SYS;
            }
        }

        // 2. one probe per GPU, all started before any is waited for
        std::vector<ProbeChild> kids((size_t)n);
        for (int i = 0; i < n && rc == GMP_OK; ++i) {
            std::vector<std::string> args = {"gpumounter-probe", "--uuid", std::string(gpus[i].uuid) + ":" + std::to_string(gpus[i].minor_number)};
            for (int k = 0; k < n_extra; ++k) {
                // a narrowed child sees its own GPU and nothing else: what the pod held before is the whole-pod step's to check
                if (strcmp(extra_args[k], "--also-mounted") == 0) {
                    ++k;
                    continue;
                }
                args.push_back(extra_args[k]);
            }
            args.push_back("--skip-p2p");
            if (not_narrowed) args.push_back("--host-view");  // more devices visible than named is then expected, not an error
            rc = spawn_probe(ctx, dirs[(size_t)i] + "/cgroup.procs", probe_exe, args, &kids[(size_t)i]);
        }
        bool timed_out = false;
        collect_probes(kids, timeout_ms, &timed_out);  // also reaps what did start when a later spawn failed
        remove_dirs();
        if (rc != GMP_OK) return rc;
        if (timed_out) {
            set_error("in-pod probes did not finish within %u ms and were killed", timeout_ms);
            return GMP_ERR_TIMEOUT;
        }
        int status = GMP_OK;
        std::string detail;
        std::vector<std::string> lines((size_t)n);
        unsigned long long start_ns = ~0ull, usable_ns = 0;
        bool all_usable = true;
        for (int i = 0; i < n; ++i) {
            int code = -1;
            const int v = probe_verdict(kids[(size_t)i], &lines[(size_t)i], &code);
            if (v != GMP_OK && status == GMP_OK) {
                status = v;
                detail = std::string(gpus[i].uuid) + ": " + (lines[(size_t)i].empty() ? gmp_last_error() : gmp_strerror(v));
                if (exit_status) *exit_status = code;
            }
            if (lines[(size_t)i].empty() || lines[(size_t)i][0] != '{') {
                lines[(size_t)i] = "{}";
                all_usable = false;
                continue;
            }
            const unsigned long long s0 = json_u64(lines[(size_t)i], "mono_start_ns"), u = json_u64(lines[(size_t)i], "mono_usable_ns");
            if (s0 != 0 && s0 < start_ns) start_ns = s0;
            if (u == 0) all_usable = false;
            if (u > usable_ns) usable_ns = u;
        }
        if (!all_usable) usable_ns = 0;
        if (start_ns == ~0ull) start_ns = 0;

        // 3. the pod as ONE process sees it: CUDA enumerates exactly the mounted set, and the P2P step across it.  After
        //    the usable instant, so its eight-GPU cuInit does not slow the per-GPU probes down.
        std::string whole = "null";
        if (status == GMP_OK) {
            std::vector<std::string> args = {"gpumounter-probe"};
            for (int i = 0; i < n; ++i) {
                args.push_back("--uuid");
                args.push_back(std::string(gpus[i].uuid) + ":" + std::to_string(gpus[i].minor_number));
            }
            for (int k = 0; k < n_extra; ++k) args.push_back(extra_args[k]);
            args.push_back("--skip-hbm");
            std::vector<ProbeChild> one(1);
            rc = spawn_probe(ctx, ctx.cgroup_dir + "/cgroup.procs", probe_exe, args, &one[0]);
            if (rc != GMP_OK) return rc;
            collect_probes(one, timeout_ms, &timed_out);
            if (timed_out) {
                set_error("whole-pod probe did not finish within %u ms and was killed", timeout_ms);
                return GMP_ERR_TIMEOUT;
            }
            std::string line;
            int code = -1;
            const int v = probe_verdict(one[0], &line, &code);
            if (!line.empty() && line[0] == '{') whole = line;
            if (v != GMP_OK) {
                status = v;
                detail = std::string("whole-pod view: ") + (whole == "null" ? gmp_last_error() : gmp_strerror(v));
                if (exit_status) *exit_status = code;
            }
        }
        if (status == GMP_OK && exit_status) *exit_status = 0;

        std::string out = "{\"status\":" + std::to_string(status) + ",\"verdict\":" + json_quote(gmp_strerror(status)) +
                          ",\"detail\":" + json_quote(detail.c_str()) + ",\"fanout\":true,\"n_gpus\":" + std::to_string(n) +
                          ",\"mono_start_ns\":" + std::to_string(start_ns) + ",\"mono_usable_ns\":" + std::to_string(usable_ns) +
                          ",\"reports\":[";
        for (int i = 0; i < n; ++i) out += (i ? "," : "") + lines[(size_t)i];
        out += "],\"whole_pod\":" + whole + "}";
        rc = emit(out, json, cap);
        if (rc != GMP_OK) return rc;
        if (status != GMP_OK) set_error("%s", detail.c_str());
        return status;
    });
}
