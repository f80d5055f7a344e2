/*
 * oracle/gmo_reference.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the pure logic and of the fork/exec mount sequence of
 * jason-gideon/GPUMounter's worker (Go), function for function, each citing the
 * reference file:line it follows.  The reference cannot be built in this image
 * (no Go toolchain, un-vendored modules, no network), so this restatement is the
 * oracle; it is used only by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs.  Nothing under gpumounter_b200/ calls it.
 *
 * PARITY PINNING.  The reference's own tests hold no golden vectors;
 * the single known answer in its tree is the comment at
 * pkg/util/cgroup/cgroup.go:47-49, checked in tests/test_oracle_reference.py.
 * Everything else is pinned only by vectors derived by hand from the cited code
 * (tests/golden/appendix_a.json): PARITY UNPINNED beyond that comment vector.
 * Third-party arithmetic restated from its published source because the
 * dependency is not vendored: runc v1.0.0-rc92 libcontainer/cgroups/systemd
 * ExpandSlice (go.mod:8; call site cgroup.go:62) and the comparison/addition of
 * k8s.io/apimachinery v0.18.6 resource.Quantity (go.mod:12; call sites
 * cgroup.go:188-211,226).
 */
#ifndef GMO_REFERENCE_H_
#define GMO_REFERENCE_H_
#include <stddef.h>
#include <stdint.h>

/* ---- pkg/device/nvidia.go ------------------------------------------------------ */
#define GMO_NVIDIA_MAJOR 195            /* nvidia.go:37 DEFAULT_NVIDA_MAJOR_NUMBER */
#define GMO_CGROUP_PERMISSION "rw"      /* nvidia.go:38 */
#define GMO_DEVICE_FILE_PERMISSION "666" /* nvidia.go:39 */
#define GMO_DEVICE_FILE_PREFIX "/dev/nvidia" /* nvidia.go:40 */
#define GMO_GPU_FREE_STATE "GPU_FREE_STATE"           /* nvidia.go:21 */
#define GMO_GPU_ALLOCATED_STATE "GPU_ALLOCATED_STATE" /* nvidia.go:22 */

typedef struct gmo_gpu { /* nvidia.go:10-17, field order kept */
    int  minor_number;
    char device_file_path[64];
    char uuid[96];
    char state[24];
    char pod_name[256];
    char ns[256];
} gmo_gpu;

void gmo_device_new(int minor, const char* uuid, gmo_gpu* out); /* nvidia.go:25-34 */
int  gmo_device_string(const gmo_gpu* g, char* buf, size_t cap); /* nvidia.go:43-50; returns length or -1 */
void gmo_device_reset_state(gmo_gpu* g);                         /* nvidia.go:52-56 */

/* ---- pkg/util/gpu/types.go ------------------------------------------------------ */
enum { GMO_MOUNT_ENTIRE = 0, GMO_MOUNT_SINGLE = 1, GMO_MOUNT_NONE = 2, GMO_MOUNT_UNKNOWN = 3 }; /* types.go:22-27 */
const char* gmo_mount_type_string(int t);
#define GMO_GPU_POOL_NAMESPACE "gpu-pool"     /* types.go:18 */
#define GMO_NVIDIA_RESOURCE_NAME "nvidia.com/gpu" /* types.go:10 */
#define GMO_KUBELET_SOCKET "/var/lib/kubelet/pod-resources/kubelet.sock" /* types.go:6-7 */

/* ---- pod model (the fields of corev1.Pod the path reads) -------------------------- */
typedef struct gmo_container_resources { /* NULL = key absent from the map; only cpu/memory count, cgroup.go:171-175 */
    const char* cpu_request;
    const char* memory_request;
    const char* cpu_limit;
    const char* memory_limit;
} gmo_container_resources;

typedef struct gmo_pod {
    const char* name;
    const char* ns;
    const char* uid;
    const char* node_name;
    const char* container_id; /* Status.ContainerStatuses[0].ContainerID, e.g. "docker://<64 hex>" */
    int n_containers;
    const gmo_container_resources* containers;
} gmo_pod;

/* ---- pkg/util/cgroup/cgroup.go ----------------------------------------------------- */
enum { GMO_QOS_GUARANTEED = 0, GMO_QOS_BURSTABLE = 1, GMO_QOS_BESTEFFORT = 2 };
enum { GMO_E_OK = 0, GMO_E_PANIC_NAME = -1, GMO_E_DRIVER = -2, GMO_E_SLICE = -3, GMO_E_BUF = -4, GMO_E_QUANTITY = -5,
       GMO_E_TARGET = -6, GMO_E_IO = -7, GMO_E_EXEC = -8, GMO_E_PANIC_INDEX = -9, GMO_E_BUSY = -10, GMO_E_ATOI = -11 };

int gmo_quantity_cmp(const char* a, const char* b, int* out); /* resource.Quantity.Cmp on the canonical serialisations */
int gmo_pod_qos(const gmo_pod* pod);                           /* cgroup.go:177-237; <0 on unparsable quantity */
const char* gmo_qos_string(int qos);                           /* "Guaranteed" / "Burstable" / "BestEffort" */

int gmo_new_cgroup_name_check(const char* component);          /* cgroup.go:32-45: GMO_E_PANIC_NAME where Go panics */
int gmo_expand_slice(const char* slice, char* out, size_t cap); /* runc rc92 systemd.ExpandSlice */
int gmo_to_systemd(const char* const* parts, int n, char* out, size_t cap);  /* cgroup.go:52-72 */
int gmo_to_cgroupfs(const char* const* parts, int n, char* out, size_t cap); /* cgroup.go:74-76 */
int gmo_get_cgroup_driver(const char* env_value, char* out, size_t cap, char* err, size_t err_cap); /* cgroup.go:78-84 */
int gmo_get_cgroup_name(const char* driver, const gmo_pod* pod, const char* container_id, char* out, size_t cap); /* cgroup.go:86-113 */
int gmo_device_group_path(const char* devices_root, const char* cgroup_path, char* out, size_t cap); /* cgroup.go:115-118 */
/* cgroup.go:120-141: lines of cgroup.procs that END in '\n'; pids is a caller array of char[16]. Returns count or <0. */
int gmo_get_cgroup_pids(const char* devices_root, const char* cgroup_path, char (*pids)[16], int cap);
/* cgroup.go:145 / :159: the shell command string. allow != 0 -> devices.allow, else devices.deny */
int gmo_permission_command(const char* devices_root, const char* cgroup_path, int minor, int allow, char* out, size_t cap);

/* ---- pkg/util/namespace/namespace.go ------------------------------------------------- */
typedef struct gmo_nsenter_config { /* namespace.go:16-39, same fields */
    int cgroup; const char* cgroup_file;
    int follow_context;
    int gid;
    int ipc; const char* ipc_file;
    int mount; const char* mount_file;
    int net; const char* net_file;
    int no_fork;
    int pid; const char* pid_file;
    int preserve_credentials;
    const char* root_directory;
    int target;
    int uid;
    int user; const char* user_file;
    int uts; const char* uts_file;
    const char* working_directory;
} gmo_nsenter_config;

/* namespace.go:70-165 then :59-60: argv[0]="nsenter", flags, program, args. Strings are written into `pool`.
 * Returns argc, or GMO_E_TARGET ("Target must be specified "), or GMO_E_BUF. */
int gmo_nsenter_argv(const gmo_nsenter_config* c, const char* program, const char* const* args, int nargs,
                     char** argv, int argv_cap, char* pool, size_t pool_cap);
int gmo_mknod_command(const char* dev_prefix, int minor, char* out, size_t cap);  /* namespace.go:168 */
int gmo_rm_command(const char* dev_prefix, int minor, char* out, size_t cap);     /* namespace.go:180 */
int gmo_kill_command(const char* const* pids, int n, char* out, size_t cap);      /* namespace.go:192 */

/* ---- pkg/util/util.go --------------------------------------------------------------- */
int gmo_contain_string(const char* const* list, int n, const char* aim);            /* util.go:198-205 */
int gmo_can_mount(int mount_type, int request_is_entire_mount);                      /* util.go:207-226 */
/* util.go:184-195: gpu process pids (NVML order: graphics then compute) that are in the cgroup's pid list.
 * out gets indices into gpu_pids; returns count (0 == Go's nil). */
int gmo_pod_gpu_processes(const char* const* cgroup_pids, int n_cgroup, const unsigned* gpu_pids, int n_gpu, int* out_idx);

/* ---- pkg/util/gpu/collector/collector.go ------------------------------------------- */
typedef struct gmo_container_devices { const char* resource_name; int n_ids; const char* const* device_ids; } gmo_container_devices;
typedef struct gmo_container_res_entry { const char* name; int n_devices; const gmo_container_devices* devices; } gmo_container_res_entry;
typedef struct gmo_pod_resources { const char* name; const char* ns; int n_containers; const gmo_container_res_entry* containers; } gmo_pod_resources;

/* collector.go:81-88; returns index or -1 */
int gmo_get_gpu_by_uuid(const gmo_gpu* list, int n, const char* uuid);
/* collector.go:113-135: reset all, then mark allocations; returns 0, or -1 on an unknown UUID (state left as the
 * reference leaves it: already-processed devices stay marked). */
int gmo_update_gpu_status(gmo_gpu* list, int n, const gmo_pod_resources* pods, int n_pods);
/* collector.go:155-161: indices of the pod's GPUs (own + slave pods', substring match). Returns count. */
int gmo_pod_gpu_resources(const gmo_gpu* list, int n, const char* pod_name, const char* ns, int* out_idx);

/* ---- pkg/util/gpu/allocator/allocator.go --------------------------------------------- */
int gmo_get_mount_type(const gmo_gpu* list, int n, const char* pod_name, const char* ns);            /* :158-187 (without the RPC-failure branch) */
/* :101-126; out_idx gets list indices; returns count (0 when emptied by the length check) */
int gmo_get_remove_gpu(const gmo_gpu* list, int n, const char* owner_name, const char* owner_ns,
                       const char* const* uuids, int n_uuids, int* out_idx);
int gmo_slave_pod_name(const char* owner, const unsigned char rand3[3], char* out, size_t cap);     /* :191-199 */
/* server.go:61-65 + allocator.go:49: number of slave pods created; -1 where Go panics (integer divide by zero). */
int gmo_slave_pod_count(int gpu_num, int is_entire_mount, int* gpu_num_per_pod);

/* ---- cmd/GPUMounter-master/main.go: gRPC result enum -> HTTP status ---------------- */
enum { GMO_ADD_SUCCESS = 0, GMO_ADD_INSUFFICIENT_GPU = 1, GMO_ADD_POD_NOT_FOUND = 2 };          /* api.proto:12-18 */
enum { GMO_REMOVE_SUCCESS = 0, GMO_REMOVE_GPU_BUSY = 1, GMO_REMOVE_POD_NOT_FOUND = 2, GMO_REMOVE_GPU_NOT_FOUND = 4 }; /* api.proto:32-39 */
/* HTTP status and response body the master writes for a worker result. http.Error appends "\n" to its message.
 * A result with no case in the switch leaves the handler with Go's implicit 200 and an empty body. */
int gmo_add_result_http(int result, const char* pod_name, const char* node_name, char* body, size_t cap);    /* main.go:103-116 */
int gmo_remove_result_http(int result, const char* pod_name, const char* node_name, const char* const* uuids, int n_uuids,
                           char* body, size_t cap); /* main.go:206-224 */

/* ---- the effectful mount path, fork/exec for fork/exec (the timed CPU baseline) ------- */
typedef struct gmo_mount_env {
    const char* devices_root;   /* reference constant "/sys/fs/cgroup/devices" (cgroup.go:116); injectable for the harness */
    const char* dev_prefix;     /* reference constant "/dev/nvidia" (nvidia.go:40); injectable for the harness */
    const char* cgroup_driver;  /* value of $CGROUP_DRIVER (cgroup.go:79) */
} gmo_mount_env;

/* util.go:17-71. Returns GMO_E_*; err receives a message. */
int gmo_mount_gpu(const gmo_mount_env* env, const gmo_pod* pod, const gmo_gpu* gpu, char* err, size_t err_cap);
/* util.go:73-147; gpu_pids = what NvidiaGPU.GetRunningProcess would return (nvidia.go:58-87). */
int gmo_unmount_gpu(const gmo_mount_env* env, const gmo_pod* pod, const gmo_gpu* gpu, int force,
                    const unsigned* gpu_pids, int n_gpu_pids, char* err, size_t err_cap);

#endif
