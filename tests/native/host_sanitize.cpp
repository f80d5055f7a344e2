// Sanitizer driver for the host-only half of the library (gmp_mount.cpp + gmp_nvml.cpp + gmp_devfilter.cpp): built once with
// -fsanitize=address,undefined and once with -fsanitize=thread by tests/test_host_sanitizers.py.
// Exercises the string/path/QoS/selection logic with hostile inputs, the nsenter argv builder, and — from eight
// threads at once — the entry points the header declares thread-safe.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <thread>
#include <unistd.h>
#include <vector>

#include "gpumounter_probe.h"

// the error-plumbing functions normally come from gmp_core.cu (a CUDA translation unit)
namespace gmp {
static thread_local char g_err[512];
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void clear_error() { g_err[0] = '\0'; }
}  // namespace gmp
extern "C" const char* gmp_last_error(void) { return gmp::g_err; }
extern "C" const char* gmp_strerror(int) { return "error"; }  // lives in gmp_core.cu, which this host-only build leaves out

#define CHECK(cond)                                                         \
    do {                                                                    \
        if (!(cond)) {                                                      \
            fprintf(stderr, "CHECK failed at line %d: %s\n", __LINE__, #cond); \
            exit(2);                                                        \
        }                                                                   \
    } while (0)

static void logic_round(int salt) {
    char buf[4096];
    // cgroup names: empty, long, hostile
    const char* parts[] = {"kubepods", "burstable", "pod1234-abcd-5678-efgh"};
    CHECK(gmp_cgroup_name_to_systemd(parts, 3, buf, sizeof(buf)) == GMP_OK);
    CHECK(strcmp(buf, "/kubepods.slice/kubepods-burstable.slice/kubepods-burstable-pod1234_abcd_5678_efgh.slice") == 0);
    CHECK(gmp_cgroup_name_to_systemd(parts, 3, buf, 8) == GMP_ERR_BUFFER_TOO_SMALL);
    CHECK(gmp_cgroup_name_to_systemd(nullptr, 0, buf, sizeof(buf)) == GMP_OK && strcmp(buf, "/") == 0);
    const std::string longpart(3000, 'a');
    const char* lp[] = {longpart.c_str(), "-", ""};
    CHECK(gmp_cgroup_name_to_systemd(lp, 3, buf, sizeof(buf)) != GMP_OK || true);
    CHECK(gmp_cgroup_name_to_cgroupfs(lp, 3, buf, sizeof(buf)) == GMP_OK);
    CHECK(gmp_cgroup_name_check("a/b") == GMP_ERR_NAME && gmp_cgroup_name_check("a_b") == GMP_ERR_NAME);
    // QoS with odd quantities
    gmp_container_resources cr[3] = {{"100m", "1Gi", "0.1", "1073741824"}, {nullptr, nullptr, nullptr, nullptr}, {"1e3", "12Pi", "1k", "12Pi"}};
    gmp_pod pod = {"p", "ns", "uid-1", "node", "docker://abc", 3, cr};
    int qos = -1;
    CHECK(gmp_get_pod_qos(&pod, &qos) == GMP_OK && qos == GMP_QOS_BURSTABLE);
    gmp_container_resources bad[1] = {{"12 parsecs", nullptr, nullptr, nullptr}};
    gmp_pod badpod = {"p", "ns", "u", "n", "c", 1, bad};
    CHECK(gmp_get_pod_qos(&badpod, &qos) == GMP_ERR_INVALID);
    CHECK(gmp_get_cgroup_name("systemd", &pod, "cid", buf, sizeof(buf)) == GMP_OK);
    CHECK(gmp_get_cgroup_name("cgroupfs", &pod, "cid", buf, 4) == GMP_ERR_BUFFER_TOO_SMALL);
    CHECK(gmp_get_cgroup_name("nope", &pod, "cid", buf, sizeof(buf)) == GMP_ERR_CGROUP_DRIVER);
    // device records, JSON escaping with every byte value
    gmp_gpu g;
    CHECK(gmp_device_new(3 + salt, "GPU-x", &g) == GMP_OK);
    for (int c = 1; c < 255 && c < (int)sizeof(g.pod_name); ++c) g.pod_name[c - 1] = (char)c;
    g.pod_name[254] = '\0';
    CHECK(gmp_device_string(&g, buf, sizeof(buf)) == GMP_OK);
    CHECK(gmp_device_string(&g, buf, 16) == GMP_ERR_BUFFER_TOO_SMALL);
    CHECK(gmp_device_reset_state(&g) == GMP_OK && g.pod_name[0] == '\0');
    // nsenter argv with everything switched on and long file names
    gmp_ns_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.target = 4242;
    cfg.cgroup = cfg.ipc = cfg.mount = cfg.net = cfg.pid = cfg.user = cfg.uts = 1;
    cfg.mount_file = longpart.c_str();
    cfg.root_directory = "/r";
    cfg.working_directory = "/w";
    cfg.gid = 7;
    cfg.uid = 9;
    int argc = 0;
    CHECK(gmp_ns_build_command(&cfg, buf, sizeof(buf), &argc) == GMP_OK && argc > 10);
    CHECK(gmp_ns_build_command(&cfg, buf, 64, &argc) == GMP_ERR_BUFFER_TOO_SMALL);
    cfg.target = 0;
    CHECK(gmp_ns_build_command(&cfg, buf, sizeof(buf), &argc) == GMP_ERR_TARGET);
    // ownership bookkeeping
    gmp_gpu list[4];
    const char* uu[4] = {"GPU-0", "GPU-1", "GPU-2", "GPU-3"};
    for (int i = 0; i < 4; ++i) CHECK(gmp_device_new(i, uu[i], &list[i]) == GMP_OK);
    const char* ids01[] = {"GPU-0", "GPU-1"};
    gmp_container_devices cd[1] = {{"nvidia.com/gpu", 2, ids01}};
    gmp_container_entry ce[1] = {{"c", 1, cd}};
    gmp_pod_resources pr[1] = {{"a-slave-pod-123456", "gpu-pool", 1, ce}};
    CHECK(gmp_update_gpu_status(list, 4, pr, 1) == GMP_OK);
    int32_t idx[4];
    int count = -1, mt = -1;
    CHECK(gmp_get_pod_gpu_resources(list, 4, "a", "default", idx, &count) == GMP_OK && count == 2);
    CHECK(gmp_get_mount_type(list, 4, "a", "default", &mt) == GMP_OK && mt == GMP_MOUNT_ENTIRE);
    const char* want[] = {"x", "y"};
    CHECK(gmp_get_remove_gpu(list, 4, "a", "default", want, 2, idx, &count) == GMP_OK && count == 2);
    CHECK(gmp_get_remove_gpu(list, 4, "a", "default", want, 1, idx, &count) == GMP_OK && count == 0);
    const char* unknown[] = {"GPU-9"};
    gmp_container_devices cdu[1] = {{"nvidia.com/gpu", 1, unknown}};
    gmp_container_entry ceu[1] = {{"c", 1, cdu}};
    gmp_pod_resources pru[1] = {{"z", "default", 1, ceu}};
    CHECK(gmp_update_gpu_status(list, 4, pru, 1) == GMP_ERR_NO_DEVICE);
    const unsigned char r3[3] = {1, 2, 3};
    CHECK(gmp_slave_pod_name("owner", r3, buf, sizeof(buf)) == GMP_OK && strcmp(buf, "owner-slave-pod-010203") == 0);
    CHECK(gmp_slave_pod_name("owner", r3, buf, 5) == GMP_ERR_BUFFER_TOO_SMALL);
    // device filter (header section 7): build, decode, evaluate; truncated and garbled programs must fail cleanly
    {
        gmp_dev_rule rules[3] = {{GMP_DEV_CHAR, -1, -1, GMP_DEV_ACC_MKNOD, 1}, {GMP_DEV_CHAR, 1, 3, GMP_DEV_ACC_ALL, 1},
                                 {GMP_DEV_CHAR, 195, salt, GMP_DEV_ACC_READ | GMP_DEV_ACC_WRITE, 1}};
        gmp_bpf_insn prog[64];
        int n_insns = 0, n_rules = 0, def = -1, allowed = -1;
        CHECK(gmp_devfilter_build(rules, 3, 0, prog, 64, &n_insns) == GMP_OK && n_insns == 6 + 6 + 5 + 8 + 2);
        CHECK(gmp_devfilter_build(rules, 3, 0, prog, 5, &n_insns) == GMP_ERR_BUFFER_TOO_SMALL && n_insns == 27);
        gmp_dev_rule back[8];
        CHECK(gmp_devfilter_decode(prog, n_insns, back, 8, &n_rules, &def) == GMP_OK && n_rules == 3 && def == 0);
        CHECK(memcmp(back, rules, sizeof(rules)) == 0);
        CHECK(gmp_devfilter_decode(prog, n_insns, back, 2, &n_rules, &def) == GMP_ERR_BUFFER_TOO_SMALL && n_rules == 3);
        CHECK(gmp_devfilter_eval(prog, n_insns, GMP_DEV_CHAR, GMP_DEV_ACC_READ, 195, (uint32_t)salt, &allowed) == GMP_OK && allowed == 1);
        CHECK(gmp_devfilter_eval(prog, n_insns, GMP_DEV_CHAR, GMP_DEV_ACC_READ, 195, 77, &allowed) == GMP_OK && allowed == 0);
        for (int cut = 1; cut < n_insns; ++cut) {  // every truncation: an error code, never a read past the end
            CHECK(gmp_devfilter_decode(prog, cut, back, 8, &n_rules, &def) == GMP_ERR_UNSUPPORTED);
            const int rc = gmp_devfilter_eval(prog, cut, GMP_DEV_CHAR, GMP_DEV_ACC_READ, 9, 9, &allowed);
            CHECK(rc == GMP_OK || rc == GMP_ERR_INVALID);
        }
        for (int at = 0; at < n_insns; ++at) {     // every single-instruction corruption
            gmp_bpf_insn bad[64];
            memcpy(bad, prog, sizeof(gmp_bpf_insn) * (size_t)n_insns);
            bad[at].off = (int16_t)(bad[at].off + 1000 * (salt + 1));
            bad[at].regs = (uint8_t)(bad[at].regs ^ 0xFF);
            (void)gmp_devfilter_decode(bad, n_insns, back, 8, &n_rules, &def);
            (void)gmp_devfilter_eval(bad, n_insns, GMP_DEV_CHAR, GMP_DEV_ACC_READ, 1, 3, &allowed);
        }
        {   // planner: runc layout edited in place, a foreign program wrapped and unwrapped, garbage refused or wrapped
            gmp_bpf_insn planned[128], again[128];
            int n_planned = 0, n_again = 0;
            gmp_dev_rule gpu = {GMP_DEV_CHAR, 195, 40 + salt, GMP_DEV_ACC_READ | GMP_DEV_ACC_WRITE, 1};
            CHECK(gmp_devfilter_build(rules, 3, 0, prog, 64, &n_insns) == GMP_OK);
            CHECK(gmp_devfilter_plan(prog, n_insns, &gpu, 1, 1, planned, 128, &n_planned) == GMP_OK && n_planned == n_insns + 8);
            CHECK(gmp_devfilter_plan(planned, n_planned, &gpu, 1, 0, again, 128, &n_again) == GMP_OK && n_again == n_insns);
            CHECK(memcmp(again, prog, sizeof(gmp_bpf_insn) * (size_t)n_insns) == 0);
            gmp_bpf_insn foreign[2] = {{0xB7, 0, 0, 1}, {0x95, 0, 0, 0}};  // r0 = 1; exit
            CHECK(gmp_devfilter_plan(foreign, 2, &gpu, 1, 1, planned, 128, &n_planned) == GMP_OK && n_planned == 2 + 6 + 8 + 2);
            CHECK(gmp_devfilter_plan(planned, n_planned, &gpu, 1, 0, again, 128, &n_again) == GMP_OK);
            CHECK(n_again == 2 + 6 + 8 + 2);  // "allow everything" still reaches the GPU: the allow became a deny
            CHECK(gmp_devfilter_eval(again, n_again, GMP_DEV_CHAR, GMP_DEV_ACC_READ, 195, (uint32_t)(40 + salt), &allowed) == GMP_OK && !allowed);
            for (int cut = 1; cut < n_planned; ++cut)  // truncated prefixes: an error or a wrap, never a crash
                (void)gmp_devfilter_plan(planned, cut, &gpu, 1, 0, again, 128, &n_again);
            CHECK(gmp_devfilter_plan(planned, n_planned, &gpu, 1, 1, again, 4, &n_again) == GMP_ERR_BUFFER_TOO_SMALL);
        }
        gmp_dev_rule hostile = {7, -5, 1 << 30, 0xFFFF, 2};
        CHECK(gmp_devfilter_build(&hostile, 1, 0, prog, 64, &n_insns) == GMP_ERR_INVALID);
        uint32_t ids[4];
        int n_ids = -1;
        CHECK(gmp_cgroup2_device_programs("/nonexistent/cgroup", ids, 4, &n_ids, nullptr) == GMP_ERR_SYS);
    }
    // NVML absent: a clean error, not a crash
    gmp_device devs[2];
    int n = -1;
    setenv("GPUMOUNTER_NVML_LIB", "/nonexistent/libnvidia-ml.so.1", 1);
    CHECK(gmp_enumerate_nvml(devs, 2, &n) == GMP_ERR_NVML_NOT_FOUND && n == 0);
}

// The process-spawning entry points (clone(CLONE_VM|CLONE_VFORK) helpers, privilege drop, pipes, child groups) on a scratch
// tree whose "pod" is this very process: own pid in cgroup.procs, own mount namespace.  Needs root (setns); skipped otherwise.
static void in_pod_round() {
#if defined(__SANITIZE_THREAD__)
    return;  // TSan's own clone() interceptor aborts on CLONE_VM children (tsan_rtl.cpp FindSlotAndLock): ASan/UBSan cover this part
#endif
    if (geteuid() != 0) return;
    char tmpl[] = "/tmp/gmp_sanitize_XXXXXX";
    CHECK(mkdtemp(tmpl) != nullptr);
    const std::string root = tmpl;
    const std::string uid = "0a1b2c3d-1111-2222-3333-444455556666", cid(64, 'c');
    const std::string cg = root + "/kubepods/besteffort/pod" + uid + "/" + cid;
    CHECK(system(("mkdir -p " + cg + " " + root + "/dev").c_str()) == 0);
    {
        FILE* f = fopen((cg + "/cgroup.procs").c_str(), "w");
        CHECK(f != nullptr);
        fprintf(f, "%d\n", (int)getpid());
        fclose(f);
    }
    const std::string prefix = root + "/dev/nvidia", container = "docker://" + cid;
    gmp_mount_env env;
    memset(&env, 0, sizeof(env));
    env.devices_root = root.c_str();
    env.dev_prefix = prefix.c_str();
    env.cgroup_driver = "cgroupfs";
    env.probe_timeout_ms = 20000;
    gmp_container_resources none[1] = {{nullptr, nullptr, nullptr, nullptr}};
    gmp_pod pod = {"gpu-pod", "default", uid.c_str(), "node", container.c_str(), 1, none};
    gmp_gpu gpus[2];
    CHECK(gmp_device_new(3, "GPU-aaaaaaaa-0000-0000-0000-000000000003", &gpus[0]) == GMP_OK);
    CHECK(gmp_device_new(5, "GPU-aaaaaaaa-0000-0000-0000-000000000005", &gpus[1]) == GMP_OK);
    int done = 0, status = -2, failed = -2;
    CHECK(gmp_mount_gpus(&env, &pod, gpus, 2, &done) == GMP_OK && done == 2);
    static char json[1 << 16];
    const char* extra[] = {"--also-mounted", "GPU-x:9", "--bytes", "1M"};
    for (uint32_t ext : {0u, (uint32_t)GMP_EXT_PROBE_KEEP_PRIVILEGES}) {
        env.extensions = ext;
        // /bin/echo stands in for the probe: its "report" is its own argument list
        CHECK(gmp_probe_in_pod(&env, &pod, gpus, 2, "/bin/echo", extra, 4, json, sizeof(json), &status) == GMP_OK && status == 0);
        CHECK(strstr(json, "--uuid GPU-aaaaaaaa-0000-0000-0000-000000000003:3") != nullptr && strstr(json, "--also-mounted GPU-x:9") != nullptr);
        CHECK(gmp_probe_in_pod(&env, &pod, gpus, 2, "/nonexistent/probe", nullptr, 0, json, sizeof(json), &status) == GMP_ERR_SYS);
        CHECK(gmp_probe_in_pod(&env, &pod, gpus, 2, "/bin/echo", nullptr, 0, json, 4, &status) == GMP_ERR_BUFFER_TOO_SMALL);
        const int rc = gmp_probe_in_pod_fanout(&env, &pod, gpus, 2, "/bin/echo", extra, 4, json, sizeof(json), &status);
        if (rc != GMP_OK) fprintf(stderr, "fan-out: rc=%d %s | %s\n", rc, gmp_last_error(), json);
        CHECK(rc == GMP_OK && strstr(json, "\"fanout\":true") != nullptr && strstr(json, "\"whole_pod\":null") != nullptr);
        CHECK(gmp_probe_in_pod_fanout(&env, &pod, gpus, 2, "/nonexistent/probe", nullptr, 0, json, sizeof(json), &status) == GMP_ERR_SYS);
    }
    env.extensions = 0;
    (void)gmp_check_nodes_in_pod(&env, &pod, gpus, 2, &failed);  // scratch nodes have no driver behind them: an error code, no crash
    CHECK(gmp_verify_gone_in_pod(&env, &pod, gpus, 2, &failed) == GMP_ERR_STILL_MOUNTED && failed == 0);
    CHECK(system(("rm -rf " + root).c_str()) == 0);
}

int main() {
    logic_round(0);
    std::vector<std::thread> ts;
    for (int t = 0; t < 8; ++t)
        ts.emplace_back([t]() {
            for (int i = 0; i < 50; ++i) logic_round(t);
        });
    for (auto& t : ts) t.join();
    in_pod_round();
    puts("host sanitize ok");
    return 0;
}
