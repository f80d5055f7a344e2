/* oracle/gmo_reference.c — TEST INFRASTRUCTURE (see gmo_reference.h). Plain C restatement of the
 * GPUMounter worker's pure logic and of its fork/exec mount sequence. */
#define _GNU_SOURCE
#include "gmo_reference.h"

#include <ctype.h>
#include <errno.h>
#include <fcntl.h>
#include <spawn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

extern char** environ;

/* ---- tiny bounded string builder -------------------------------------------------- */
typedef struct sb {
    char* p;
    size_t cap, len;
    int overflow;
} sb;

static void sb_init(sb* s, char* buf, size_t cap) {
    s->p = buf;
    s->cap = cap;
    s->len = 0;
    s->overflow = cap == 0;
    if (cap) buf[0] = '\0';
}

static void sb_putn(sb* s, const char* t, size_t n) {
    if (s->overflow) return;
    if (s->len + n + 1 > s->cap) {
        s->overflow = 1;
        return;
    }
    memcpy(s->p + s->len, t, n);
    s->len += n;
    s->p[s->len] = '\0';
}

static void sb_put(sb* s, const char* t) { sb_putn(s, t, strlen(t)); }

static void sb_int(sb* s, long v) { /* strconv.Itoa */
    char tmp[32];
    snprintf(tmp, sizeof(tmp), "%ld", v);
    sb_put(s, tmp);
}

static void copy_str(char* dst, size_t cap, const char* src) {
    if (cap == 0) return;
    size_t n = strlen(src);
    if (n >= cap) n = cap - 1;
    memcpy(dst, src, n);
    dst[n] = '\0';
}

/* ---- pkg/device/nvidia.go ------------------------------------------------------------ */
void gmo_device_new(int minor, const char* uuid, gmo_gpu* out) { /* nvidia.go:25-34 */
    memset(out, 0, sizeof(*out));
    out->minor_number = minor;
    snprintf(out->device_file_path, sizeof(out->device_file_path), "%s%d", GMO_DEVICE_FILE_PREFIX, minor);
    copy_str(out->uuid, sizeof(out->uuid), uuid);
    copy_str(out->state, sizeof(out->state), GMO_GPU_FREE_STATE);
    out->pod_name[0] = '\0';
    out->ns[0] = '\0';
}

void gmo_device_reset_state(gmo_gpu* g) { /* nvidia.go:52-56 */
    g->pod_name[0] = '\0';
    g->ns[0] = '\0';
    copy_str(g->state, sizeof(g->state), GMO_GPU_FREE_STATE);
}

/* Width of the UTF-8 sequence at c as Go's unicode/utf8 sees it, or 1 for an invalid one (table of RFC 3629: no
 * overlong forms, no surrogates, nothing above U+10FFFF). */
static int go_rune_width(const unsigned char* c) {
    unsigned char lo = 0x80, hi = 0xBF;
    int n;
    if (c[0] >= 0xC2 && c[0] <= 0xDF) n = 2;
    else if (c[0] >= 0xE0 && c[0] <= 0xEF) { n = 3; if (c[0] == 0xE0) lo = 0xA0; if (c[0] == 0xED) hi = 0x9F; }
    else if (c[0] >= 0xF0 && c[0] <= 0xF4) { n = 4; if (c[0] == 0xF0) lo = 0x90; if (c[0] == 0xF4) hi = 0x8F; }
    else return 1;
    if (c[1] < lo || c[1] > hi) return 1;
    for (int k = 2; k < n; ++k)
        if (c[k] < 0x80 || c[k] > 0xBF) return 1;
    return n;
}

/* encoding/json string escaping with the default HTML-safe mode that json.Marshal uses. */
static void json_string(sb* s, const char* t) {
    sb_put(s, "\"");
    for (const unsigned char* c = (const unsigned char*)t; *c; ++c) {
        char tmp[8];
        switch (*c) {
            case '"': sb_put(s, "\\\""); break;
            case '\\': sb_put(s, "\\\\"); break;
            case '\n': sb_put(s, "\\n"); break;
            case '\r': sb_put(s, "\\r"); break;
            case '\t': sb_put(s, "\\t"); break;
            case '<': case '>': case '&':
                snprintf(tmp, sizeof(tmp), "\\u%04x", *c);
                sb_put(s, tmp);
                break;
            default:
                if (*c < 0x20) {
                    snprintf(tmp, sizeof(tmp), "\\u%04x", *c);
                    sb_put(s, tmp);
                } else if (*c < 0x80) {
                    sb_putn(s, (const char*)c, 1);
                } else {
                    /* utf8.DecodeRuneInString: RuneError with width 1 is written as \ufffd; U+2028 / U+2029 are
                     * escaped (they break JavaScript); every other rune is copied (encoding/json encode.go, string()) */
                    int w = go_rune_width(c);
                    if (w == 1) {
                        sb_put(s, "\\ufffd");
                    } else if (w == 3 && c[0] == 0xE2 && c[1] == 0x80 && (c[2] == 0xA8 || c[2] == 0xA9)) {
                        sb_put(s, c[2] == 0xA8 ? "\\u2028" : "\\u2029");
                        c += 2;
                    } else {
                        sb_putn(s, (const char*)c, (size_t)w);
                        c += w - 1;
                    }
                }
        }
    }
    sb_put(s, "\"");
}

int gmo_device_string(const gmo_gpu* g, char* buf, size_t cap) { /* nvidia.go:43-50: json.Marshal, struct field order */
    sb s;
    sb_init(&s, buf, cap);
    sb_put(&s, "{\"MinorNumber\":");
    sb_int(&s, g->minor_number);
    sb_put(&s, ",\"DeviceFilePath\":");
    json_string(&s, g->device_file_path);
    sb_put(&s, ",\"UUID\":");
    json_string(&s, g->uuid);
    sb_put(&s, ",\"State\":");
    json_string(&s, g->state);
    sb_put(&s, ",\"PodName\":");
    json_string(&s, g->pod_name);
    sb_put(&s, ",\"Namespace\":");
    json_string(&s, g->ns);
    sb_put(&s, "}");
    return s.overflow ? -1 : (int)s.len;
}

const char* gmo_mount_type_string(int t) { /* types.go:22-27 */
    switch (t) {
        case GMO_MOUNT_ENTIRE: return "entire-mount";
        case GMO_MOUNT_SINGLE: return "single-mount";
        case GMO_MOUNT_NONE: return "no-mount";
        default: return "unknown-mount";
    }
}

/* ---- resource.Quantity (k8s.io/apimachinery v0.18.6), value semantics only ------------
 * Grammar: <sign><digits>[.<digits>][<suffix>] with suffix in
 *   binary SI: Ki Mi Gi Ti Pi Ei | decimal SI: n u m "" k M G T P E | decimal exponent: e<N> / E<N>.
 * Value kept exactly as mantissa * 10^exp10 * 2^exp2 in a 128-bit integer scaled to nano units. */
typedef __int128 i128;

static int quantity_parse(const char* str, i128* nano) {
    if (str == NULL) return GMO_E_QUANTITY;
    const char* p = str;
    int neg = 0;
    if (*p == '+' || *p == '-') neg = (*p++ == '-');
    i128 mant = 0;
    int digits = 0, frac = 0;
    while (isdigit((unsigned char)*p)) {
        mant = mant * 10 + (*p++ - '0');
        ++digits;
    }
    if (*p == '.') {
        ++p;
        while (isdigit((unsigned char)*p)) {
            mant = mant * 10 + (*p++ - '0');
            ++digits;
            ++frac;
        }
    }
    if (digits == 0) return GMO_E_QUANTITY;
    int exp10 = -frac;
    int exp2 = 0;
    if (p[0] && p[1] == 'i' && p[2] == '\0') {
        switch (p[0]) {
            case 'K': exp2 = 10; break;
            case 'M': exp2 = 20; break;
            case 'G': exp2 = 30; break;
            case 'T': exp2 = 40; break;
            case 'P': exp2 = 50; break;
            case 'E': exp2 = 60; break;
            default: return GMO_E_QUANTITY;
        }
    } else if ((p[0] == 'e' || p[0] == 'E') && (isdigit((unsigned char)p[1]) || ((p[1] == '-' || p[1] == '+') && isdigit((unsigned char)p[2])))) {
        char* end = NULL;
        long e = strtol(p + 1, &end, 10);
        if (*end != '\0' || e > 30 || e < -30) return GMO_E_QUANTITY;
        exp10 += (int)e;
    } else if (p[0] != '\0' && p[1] == '\0') {
        switch (p[0]) {
            case 'n': exp10 += -9; break;
            case 'u': exp10 += -6; break;
            case 'm': exp10 += -3; break;
            case 'k': exp10 += 3; break;
            case 'M': exp10 += 6; break;
            case 'G': exp10 += 9; break;
            case 'T': exp10 += 12; break;
            case 'P': exp10 += 15; break;
            case 'E': exp10 += 18; break;
            default: return GMO_E_QUANTITY;
        }
    } else if (p[0] != '\0') {
        return GMO_E_QUANTITY;
    }
    /* to nano units; Quantity itself rounds up anything finer than nano */
    exp10 += 9;
    i128 v = mant;
    for (int i = 0; i < exp2; ++i) v *= 2;
    while (exp10 > 0) {
        v *= 10;
        --exp10;
    }
    while (exp10 < 0) {
        i128 q = v / 10;
        if (q * 10 != v) q += 1; /* round up, as resource.ParseQuantity does */
        v = q;
        ++exp10;
    }
    *nano = neg ? -v : v;
    return GMO_E_OK;
}

int gmo_quantity_cmp(const char* a, const char* b, int* out) {
    i128 x, y;
    if (quantity_parse(a, &x) != 0 || quantity_parse(b, &y) != 0) return GMO_E_QUANTITY;
    *out = x < y ? -1 : (x > y ? 1 : 0);
    return GMO_E_OK;
}

/* ---- cgroup.go:177-237 GetPodQOS -------------------------------------------------------- */
const char* gmo_qos_string(int qos) {
    switch (qos) {
        case GMO_QOS_GUARANTEED: return "Guaranteed";
        case GMO_QOS_BURSTABLE: return "Burstable";
        case GMO_QOS_BESTEFFORT: return "BestEffort";
        default: return "";
    }
}

int gmo_pod_qos(const gmo_pod* pod) {
    /* requests / limits maps keyed by the two supported names: index 0 = cpu, 1 = memory (cgroup.go:171-175) */
    int have_req[2] = {0, 0}, have_lim[2] = {0, 0};
    i128 req[2] = {0, 0}, lim[2] = {0, 0};
    int is_guaranteed = 1; /* cgroup.go:181 */
    for (int c = 0; c < pod->n_containers; ++c) {
        const gmo_container_resources* r = &pod->containers[c];
        const char* rq[2] = {r->cpu_request, r->memory_request};
        const char* lm[2] = {r->cpu_limit, r->memory_limit};
        for (int k = 0; k < 2; ++k) { /* cgroup.go:184-197 */
            if (rq[k] == NULL) continue;
            i128 q;
            if (quantity_parse(rq[k], &q) != 0) return GMO_E_QUANTITY;
            if (q > 0) {
                req[k] = have_req[k] ? req[k] + q : q;
                have_req[k] = 1;
            }
        }
        int found[2] = {0, 0}; /* qosLimitsFound, cgroup.go:199 */
        for (int k = 0; k < 2; ++k) { /* cgroup.go:200-214 */
            if (lm[k] == NULL) continue;
            i128 q;
            if (quantity_parse(lm[k], &q) != 0) return GMO_E_QUANTITY;
            if (q > 0) {
                found[k] = 1;
                lim[k] = have_lim[k] ? lim[k] + q : q;
                have_lim[k] = 1;
            }
        }
        if (!(found[0] && found[1])) is_guaranteed = 0; /* cgroup.go:216-218 */
    }
    int n_req = have_req[0] + have_req[1], n_lim = have_lim[0] + have_lim[1];
    if (n_req == 0 && n_lim == 0) return GMO_QOS_BESTEFFORT; /* cgroup.go:220-222 */
    if (is_guaranteed) {                                      /* cgroup.go:224-231 */
        for (int k = 0; k < 2; ++k) {
            if (!have_req[k]) continue;
            if (!have_lim[k] || lim[k] != req[k]) {
                is_guaranteed = 0;
                break;
            }
        }
    }
    if (is_guaranteed && n_req == n_lim) return GMO_QOS_GUARANTEED; /* cgroup.go:232-235 */
    return GMO_QOS_BURSTABLE;
}

/* ---- cgroup names ----------------------------------------------------------------------- */
int gmo_new_cgroup_name_check(const char* component) { /* cgroup.go:33-40 */
    if (strchr(component, '/') != NULL || strchr(component, '_') != NULL) return GMO_E_PANIC_NAME;
    return GMO_E_OK;
}

/* runc v1.0.0-rc92 libcontainer/cgroups/systemd/common.go ExpandSlice */
int gmo_expand_slice(const char* slice, char* out, size_t cap) {
    static const char suffix[] = ".slice";
    const size_t sl = strlen(suffix), n = strlen(slice);
    /* name has to end with ".slice" */
    if (n < sl || strcmp(slice + n - sl, suffix) != 0) return GMO_E_SLICE;
    /* path separators are not allowed */
    if (strchr(slice, '/') != NULL) return GMO_E_SLICE;
    const size_t name_len = n - sl;
    sb path;
    sb_init(&path, out, cap);
    /* "-.slice" is the root slice */
    if (name_len == 1 && slice[0] == '-') {
        sb_put(&path, "/");
        return path.overflow ? GMO_E_BUF : GMO_E_OK;
    }
    char prefix[1024];
    sb pre;
    sb_init(&pre, prefix, sizeof(prefix));
    size_t start = 0;
    for (size_t i = 0; i <= name_len; ++i) {
        if (i == name_len || slice[i] == '-') {
            size_t len = i - start;
            if (len == 0) return GMO_E_SLICE; /* "a--b.slice", "-a.slice", "a-.slice", ".slice" */
            sb_put(&path, "/");
            sb_put(&path, prefix);
            sb_putn(&path, slice + start, len);
            sb_put(&path, suffix);
            sb_putn(&pre, slice + start, len);
            sb_put(&pre, "-");
            start = i + 1;
        }
    }
    return (path.overflow || pre.overflow) ? GMO_E_BUF : GMO_E_OK;
}

int gmo_to_systemd(const char* const* parts, int n, char* out, size_t cap) { /* cgroup.go:52-68 */
    if (n == 0 || (n == 1 && parts[0][0] == '\0')) {
        copy_str(out, cap, "/");
        return cap >= 2 ? GMO_E_OK : GMO_E_BUF;
    }
    char joined[2048];
    sb j;
    sb_init(&j, joined, sizeof(joined));
    for (int i = 0; i < n; ++i) {
        if (i) sb_put(&j, "-");
        for (const char* c = parts[i]; *c; ++c) { /* escapeSystemdCgroupName, cgroup.go:70-72 */
            char ch = *c == '-' ? '_' : *c;
            sb_putn(&j, &ch, 1);
        }
    }
    sb_put(&j, ".slice");
    if (j.overflow) return GMO_E_BUF;
    return gmo_expand_slice(joined, out, cap); /* a failure here is the Go panic at cgroup.go:63-66 */
}

/* path.Join semantics for elements that contain no '/': drop empties, then path.Clean */
int gmo_to_cgroupfs(const char* const* parts, int n, char* out, size_t cap) { /* cgroup.go:74-76 */
    const char* stack[64];
    int depth = 0, dotdot = 0, any = 0;
    for (int i = 0; i < n; ++i) {
        if (parts[i][0] == '\0') continue;
        any = 1;
        if (strcmp(parts[i], ".") == 0) continue;
        if (strcmp(parts[i], "..") == 0) {
            if (depth > 0) --depth;
            else ++dotdot;
            continue;
        }
        if (depth == 64) return GMO_E_BUF;
        stack[depth++] = parts[i];
    }
    sb s;
    sb_init(&s, out, cap);
    sb_put(&s, "/");
    if (any) {
        int first = 1;
        for (int i = 0; i < dotdot; ++i) {
            if (!first) sb_put(&s, "/");
            sb_put(&s, "..");
            first = 0;
        }
        for (int i = 0; i < depth; ++i) {
            if (!first) sb_put(&s, "/");
            sb_put(&s, stack[i]);
            first = 0;
        }
        if (first) sb_put(&s, "."); /* path.Clean of "." */
    }
    return s.overflow ? GMO_E_BUF : GMO_E_OK;
}

int gmo_get_cgroup_driver(const char* env_value, char* out, size_t cap, char* err, size_t err_cap) { /* cgroup.go:78-84 */
    const char* v = env_value ? env_value : ""; /* os.Getenv of an unset variable is "" */
    if (strcmp(v, "systemd") != 0 && strcmp(v, "cgroupfs") != 0) {
        if (err && err_cap) snprintf(err, err_cap, "unsupported cgroup driver: %s", v);
        if (cap) out[0] = '\0';
        return GMO_E_DRIVER;
    }
    copy_str(out, cap, v);
    return GMO_E_OK;
}

int gmo_get_cgroup_name(const char* driver, const gmo_pod* pod, const char* container_id, char* out, size_t cap) {
    /* cgroup.go:86-113 */
    const char* parts[4];
    int n = 0;
    if (gmo_new_cgroup_name_check("kubepods") != 0) return GMO_E_PANIC_NAME;
    parts[n++] = "kubepods"; /* containerRoot */
    int qos = gmo_pod_qos(pod);
    if (qos < 0) return qos;
    switch (qos) { /* cgroup.go:92-99; strings.ToLower of the QoS class */
        case GMO_QOS_GUARANTEED: break;
        case GMO_QOS_BURSTABLE: parts[n++] = "burstable"; break;
        case GMO_QOS_BESTEFFORT: parts[n++] = "besteffort"; break;
    }
    char pod_container[512];
    snprintf(pod_container, sizeof(pod_container), "pod%s", pod->uid); /* cgroup.go:101 */
    if (gmo_new_cgroup_name_check(pod_container) != 0) return GMO_E_PANIC_NAME; /* cgroup.go:102 */
    parts[n++] = pod_container;
    char base[2048];
    sb s;
    sb_init(&s, out, cap);
    if (strcmp(driver, "systemd") == 0) { /* cgroup.go:105-106 */
        int rc = gmo_to_systemd(parts, n, base, sizeof(base));
        if (rc != 0) return rc;
        sb_put(&s, base);
        sb_put(&s, "/docker-");
        sb_put(&s, container_id);
        sb_put(&s, ".scope");
    } else if (strcmp(driver, "cgroupfs") == 0) { /* cgroup.go:107-108 */
        int rc = gmo_to_cgroupfs(parts, n, base, sizeof(base));
        if (rc != 0) return rc;
        sb_put(&s, base);
        sb_put(&s, "/");
        sb_put(&s, container_id);
    } else {
        return GMO_E_DRIVER; /* cgroup.go:112 "unsupported cgroup driver" */
    }
    return s.overflow ? GMO_E_BUF : GMO_E_OK;
}

int gmo_device_group_path(const char* devices_root, const char* cgroup_path, char* out, size_t cap) { /* cgroup.go:115-118 */
    sb s;
    sb_init(&s, out, cap);
    sb_put(&s, devices_root ? devices_root : "/sys/fs/cgroup/devices");
    sb_put(&s, cgroup_path);
    return s.overflow ? GMO_E_BUF : GMO_E_OK;
}

int gmo_get_cgroup_pids(const char* devices_root, const char* cgroup_path, char (*pids)[16], int cap) { /* cgroup.go:120-141 */
    char dir[4096], file[4200];
    if (gmo_device_group_path(devices_root, cgroup_path, dir, sizeof(dir)) != 0) return GMO_E_BUF;
    snprintf(file, sizeof(file), "%s/cgroup.procs", dir);
    FILE* f = fopen(file, "r");
    if (f == NULL) return GMO_E_IO;
    int n = 0;
    char line[64];
    size_t len = 0;
    int ch;
    /* bufio.Reader.ReadString('\n'): a final fragment without '\n' comes back WITH io.EOF and is dropped (cgroup.go:132-135) */
    while ((ch = fgetc(f)) != EOF) {
        if (ch == '\n') {
            line[len] = '\0';
            if (n < cap) copy_str(pids[n], 16, line);
            ++n;
            len = 0;
        } else if (len + 1 < sizeof(line)) {
            line[len++] = (char)ch;
        }
    }
    fclose(f);
    return n;
}

int gmo_permission_command(const char* devices_root, const char* cgroup_path, int minor, int allow, char* out, size_t cap) {
    /* cgroup.go:145 and :159 */
    char dir[4096];
    if (gmo_device_group_path(devices_root, cgroup_path, dir, sizeof(dir)) != 0) return GMO_E_BUF;
    sb s;
    sb_init(&s, out, cap);
    sb_put(&s, "echo 'c ");
    sb_int(&s, GMO_NVIDIA_MAJOR);
    sb_put(&s, ":");
    sb_int(&s, minor);
    sb_put(&s, " " GMO_CGROUP_PERMISSION "' > ");
    sb_put(&s, dir);
    sb_put(&s, allow ? "/devices.allow" : "/devices.deny");
    return s.overflow ? GMO_E_BUF : GMO_E_OK;
}

/* ---- namespace.go ---------------------------------------------------------------------------- */
typedef struct argv_builder {
    char** argv;
    int cap, n;
    sb pool;
    int overflow;
} argv_builder;

static void ab_add(argv_builder* b, const char* s) {
    if (b->n + 1 >= b->cap) {
        b->overflow = 1;
        return;
    }
    size_t at = b->pool.len;
    sb_putn(&b->pool, s, strlen(s) + 1); /* keep the NUL inside the pool */
    if (b->pool.overflow) {
        b->overflow = 1;
        return;
    }
    b->argv[b->n++] = b->pool.p + at;
    b->argv[b->n] = NULL;
}

static void ab_flag(argv_builder* b, const char* with_file_prefix, const char* file, const char* bare) {
    if (file != NULL && file[0] != '\0') { /* fmt.Sprintf("--x=%s", file) */
        char tmp[4200];
        snprintf(tmp, sizeof(tmp), "%s%s", with_file_prefix, file);
        ab_add(b, tmp);
    } else {
        ab_add(b, bare);
    }
}

int gmo_nsenter_argv(const gmo_nsenter_config* c, const char* program, const char* const* args, int nargs, char** argv,
                     int argv_cap, char* pool, size_t pool_cap) {
    if (c->target == 0) return GMO_E_TARGET; /* namespace.go:71-73 */
    argv_builder b;
    b.argv = argv;
    b.cap = argv_cap;
    b.n = 0;
    b.overflow = 0;
    sb_init(&b.pool, pool, pool_cap);
    char num[32];
    ab_add(&b, "nsenter"); /* exec.CommandContext(ctx, "nsenter", args...) namespace.go:162 */
    ab_add(&b, "--target");
    snprintf(num, sizeof(num), "%d", c->target);
    ab_add(&b, num); /* namespace.go:76 */
    if (c->cgroup) ab_flag(&b, "--cgroup=", c->cgroup_file, "--cgroup");    /* :78-84 */
    if (c->follow_context) ab_add(&b, "--follow-context");                  /* :86-88 */
    if (c->gid != 0) {                                                        /* :90-92 */
        ab_add(&b, "--setgid");
        snprintf(num, sizeof(num), "%d", c->gid);
        ab_add(&b, num);
    }
    if (c->ipc) ab_flag(&b, "--ip=", c->ipc_file, "--ipc");                 /* :94-100; "--ip=" is the reference's spelling */
    if (c->mount) ab_flag(&b, "--mount=", c->mount_file, "--mount");        /* :102-108 */
    if (c->net) ab_flag(&b, "--net=", c->net_file, "--net");                /* :110-116 */
    if (c->no_fork) ab_add(&b, "--no-fork");                                /* :118-120 */
    if (c->pid) ab_flag(&b, "--pid=", c->pid_file, "--pid");                /* :122-128 */
    if (c->preserve_credentials) ab_add(&b, "--preserve-credentials");      /* :130-132 */
    if (c->root_directory && c->root_directory[0]) {                          /* :134-136 */
        ab_add(&b, "--root");
        ab_add(&b, c->root_directory);
    }
    if (c->uid != 0) { /* :138-140 */
        ab_add(&b, "--setuid");
        snprintf(num, sizeof(num), "%d", c->uid);
        ab_add(&b, num);
    }
    if (c->user) ab_flag(&b, "--user=", c->user_file, "--user"); /* :142-148 */
    if (c->uts) ab_flag(&b, "--uts=", c->uts_file, "--uts");     /* :150-156 */
    if (c->working_directory && c->working_directory[0]) {         /* :158-160 */
        ab_add(&b, "--wd");
        ab_add(&b, c->working_directory);
    }
    ab_add(&b, program); /* namespace.go:59 */
    for (int i = 0; i < nargs; ++i) ab_add(&b, args[i]); /* namespace.go:60 */
    return b.overflow ? GMO_E_BUF : b.n;
}

int gmo_mknod_command(const char* dev_prefix, int minor, char* out, size_t cap) { /* namespace.go:168 */
    sb s;
    sb_init(&s, out, cap);
    sb_put(&s, "mknod -m " GMO_DEVICE_FILE_PERMISSION " ");
    sb_put(&s, dev_prefix ? dev_prefix : GMO_DEVICE_FILE_PREFIX);
    sb_int(&s, minor);
    sb_put(&s, " c ");
    sb_int(&s, GMO_NVIDIA_MAJOR);
    sb_put(&s, " ");
    sb_int(&s, minor);
    return s.overflow ? GMO_E_BUF : GMO_E_OK;
}

int gmo_rm_command(const char* dev_prefix, int minor, char* out, size_t cap) { /* namespace.go:180 */
    sb s;
    sb_init(&s, out, cap);
    sb_put(&s, "rm ");
    sb_put(&s, dev_prefix ? dev_prefix : GMO_DEVICE_FILE_PREFIX);
    sb_int(&s, minor);
    return s.overflow ? GMO_E_BUF : GMO_E_OK;
}

int gmo_kill_command(const char* const* pids, int n, char* out, size_t cap) { /* namespace.go:192 */
    sb s;
    sb_init(&s, out, cap);
    sb_put(&s, "kill ");
    for (int i = 0; i < n; ++i) {
        if (i) sb_put(&s, " ");
        sb_put(&s, pids[i]);
    }
    return s.overflow ? GMO_E_BUF : GMO_E_OK;
}

/* ---- util.go ------------------------------------------------------------------------------------ */
int gmo_contain_string(const char* const* list, int n, const char* aim) { /* util.go:198-205 */
    for (int i = 0; i < n; ++i)
        if (strcmp(list[i], aim) == 0) return 1;
    return 0;
}

int gmo_can_mount(int mount_type, int request_is_entire_mount) { /* util.go:207-226 */
    if (mount_type == GMO_MOUNT_UNKNOWN) return 0;
    if (mount_type != GMO_MOUNT_NONE && request_is_entire_mount) return 0;
    if (mount_type == GMO_MOUNT_ENTIRE) return 0;
    return 1;
}

int gmo_pod_gpu_processes(const char* const* cgroup_pids, int n_cgroup, const unsigned* gpu_pids, int n_gpu, int* out_idx) {
    /* util.go:184-189 */
    int n = 0;
    for (int i = 0; i < n_gpu; ++i) {
        char tmp[16];
        snprintf(tmp, sizeof(tmp), "%d", (int)gpu_pids[i]); /* strconv.Itoa(int(processInfo.Pid)) */
        if (gmo_contain_string(cgroup_pids, n_cgroup, tmp)) out_idx[n++] = i;
    }
    return n;
}

/* ---- collector.go --------------------------------------------------------------------------------- */
int gmo_get_gpu_by_uuid(const gmo_gpu* list, int n, const char* uuid) { /* collector.go:81-88 */
    for (int i = 0; i < n; ++i)
        if (strcmp(list[i].uuid, uuid) == 0) return i;
    return -1;
}

int gmo_update_gpu_status(gmo_gpu* list, int n, const gmo_pod_resources* pods, int n_pods) { /* collector.go:113-135 */
    for (int i = 0; i < n; ++i) gmo_device_reset_state(&list[i]); /* resetGPUStatus, collector.go:140-144 */
    for (int p = 0; p < n_pods; ++p)
        for (int c = 0; c < pods[p].n_containers; ++c)
            for (int d = 0; d < pods[p].containers[c].n_devices; ++d) {
                const gmo_container_devices* dev = &pods[p].containers[c].devices[d];
                if (strcmp(dev->resource_name, GMO_NVIDIA_RESOURCE_NAME) != 0) continue; /* :118-120 */
                for (int u = 0; u < dev->n_ids; ++u) {
                    int idx = gmo_get_gpu_by_uuid(list, n, dev->device_ids[u]);
                    if (idx < 0) return -1; /* :123-125 */
                    copy_str(list[idx].state, sizeof(list[idx].state), GMO_GPU_ALLOCATED_STATE);
                    copy_str(list[idx].pod_name, sizeof(list[idx].pod_name), pods[p].name);
                    copy_str(list[idx].ns, sizeof(list[idx].ns), pods[p].ns);
                }
            }
    return 0;
}

int gmo_pod_gpu_resources(const gmo_gpu* list, int n, const char* pod_name, const char* ns, int* out_idx) { /* collector.go:155-161 */
    char needle[600];
    snprintf(needle, sizeof(needle), "%s-slave-pod-", pod_name);
    int k = 0;
    for (int i = 0; i < n; ++i) {
        int own = strcmp(list[i].pod_name, pod_name) == 0 && strcmp(list[i].ns, ns) == 0;
        int slave = strstr(list[i].pod_name, needle) != NULL && strcmp(list[i].ns, GMO_GPU_POOL_NAMESPACE) == 0;
        if (own || slave) out_idx[k++] = i;
    }
    return k;
}

/* ---- allocator.go ------------------------------------------------------------------------------------ */
int gmo_get_mount_type(const gmo_gpu* list, int n, const char* pod_name, const char* ns) { /* allocator.go:158-187 */
    int idx[64];
    if (n > 64) n = 64;
    int k = gmo_pod_gpu_resources(list, n, pod_name, ns, idx);
    if (k == 0) return GMO_MOUNT_NONE; /* :167-169 */
    int distinct = 0, gpu_num = 0;
    for (int i = 0; i < k; ++i) { /* :171-178: set of slave pod names */
        const char* owner = list[idx[i]].pod_name;
        if (strcmp(owner, pod_name) != 0) {
            int seen = 0;
            for (int j = 0; j < i; ++j)
                if (strcmp(list[idx[j]].pod_name, pod_name) != 0 && strcmp(list[idx[j]].pod_name, owner) == 0) seen = 1;
            if (!seen) ++distinct;
        }
        ++gpu_num;
    }
    return distinct < gpu_num ? GMO_MOUNT_ENTIRE : GMO_MOUNT_SINGLE; /* :182-186 */
}

int gmo_get_remove_gpu(const gmo_gpu* list, int n, const char* owner_name, const char* owner_ns, const char* const* uuids,
                       int n_uuids, int* out_idx) { /* allocator.go:101-126 */
    int idx[64];
    if (n > 64) n = 64;
    int k = gmo_pod_gpu_resources(list, n, owner_name, owner_ns, idx);
    int mount_type = gmo_get_mount_type(list, n, owner_name, owner_ns);
    int m = 0;
    for (int i = 0; i < k; ++i) { /* :112-119 */
        const gmo_gpu* g = &list[idx[i]];
        if ((mount_type == GMO_MOUNT_ENTIRE || gmo_contain_string(uuids, n_uuids, g->uuid)) && strcmp(g->pod_name, owner_name) != 0)
            out_idx[m++] = idx[i];
    }
    if (n_uuids != m) return 0; /* :121-123 "if exists unmatch gpu, return empty" */
    return m;
}

int gmo_slave_pod_name(const char* owner, const unsigned char rand3[3], char* out, size_t cap) { /* allocator.go:191-199 */
    int w = snprintf(out, cap, "%s-slave-pod-%02x%02x%02x", owner, rand3[0], rand3[1], rand3[2]); /* fmt.Sprintf("%x", []byte) */
    return (w < 0 || (size_t)w >= cap) ? GMO_E_BUF : GMO_E_OK;
}

int gmo_slave_pod_count(int gpu_num, int is_entire_mount, int* gpu_num_per_pod) { /* server.go:61-65, allocator.go:49 */
    int per = 1;
    if (is_entire_mount) per = gpu_num;
    if (gpu_num_per_pod) *gpu_num_per_pod = per;
    if (per == 0) return -1; /* Go: runtime panic, integer divide by zero */
    int count = gpu_num / per; /* Go's / truncates toward zero, like C */
    return count < 0 ? 0 : count; /* the for loop runs zero times for a negative bound */
}

/* ---- master: enum -> HTTP ------------------------------------------------------------------------------ */
int gmo_add_result_http(int result, const char* pod_name, const char* node_name, char* body, size_t cap) {
    /* cmd/GPUMounter-master/main.go:103-116 */
    sb s;
    sb_init(&s, body, cap);
    switch (result) {
        case GMO_ADD_SUCCESS:
            sb_put(&s, "Add GPU Success\n"); /* :106 */
            return 200;
        case GMO_ADD_INSUFFICIENT_GPU:
            sb_put(&s, "Insufficient GPU on Node: "); /* :110 */
            sb_put(&s, node_name);
            sb_put(&s, "\n");
            return 500;
        case GMO_ADD_POD_NOT_FOUND:
            sb_put(&s, "No Pod"); /* :114, no space in the reference */
            sb_put(&s, pod_name);
            sb_put(&s, " on Node: ");
            sb_put(&s, node_name);
            sb_put(&s, "\n");
            return 400;
        default:
            return 200;
    }
}

int gmo_remove_result_http(int result, const char* pod_name, const char* node_name, const char* const* uuids, int n_uuids,
                           char* body, size_t cap) { /* main.go:206-224 */
    sb s;
    sb_init(&s, body, cap);
    switch (result) {
        case GMO_REMOVE_POD_NOT_FOUND: /* :207-211 */
            sb_put(&s, "No Pod");
            sb_put(&s, pod_name);
            sb_put(&s, " on Node: ");
            sb_put(&s, node_name);
            sb_put(&s, "\n");
            return 400;
        case GMO_REMOVE_GPU_BUSY: /* :212-215 */
            sb_put(&s, "Pod: ");
            sb_put(&s, pod_name);
            sb_put(&s, " has running processes on GPU: ");
            for (int i = 0; i < n_uuids; ++i) {
                if (i) sb_put(&s, ", ");
                sb_put(&s, uuids[i]);
            }
            sb_put(&s, "\n");
            return 400;
        case GMO_REMOVE_GPU_NOT_FOUND: /* :216-219 */
            sb_put(&s, "Invalid UUIDs: ");
            for (int i = 0; i < n_uuids; ++i) {
                if (i) sb_put(&s, ", ");
                sb_put(&s, uuids[i]);
            }
            sb_put(&s, "\n");
            return 400;
        case GMO_REMOVE_SUCCESS: /* :220-223 */
            sb_put(&s, "Remove GPU Success\n");
            return 200;
        default:
            return 200;
    }
}

/* ---- fork/exec helpers (os/exec semantics) ---------------------------------------------------------------- */
/* Runs argv via PATH lookup, captures stdout+stderr together, returns the exit status or -1.
 * Go's os/exec starts children with clone(CLONE_VFORK|CLONE_VM) (syscall.forkExec), whose cost does not grow with
 * the parent's address space; posix_spawnp is the same mechanism in C, so the timed baseline is not inflated by the
 * size of the process that hosts it. */
static int run_capture(char* const argv[], char* out, size_t out_cap) {
    int pfd[2];
    if (pipe(pfd) != 0) return -1;
    posix_spawn_file_actions_t fa;
    posix_spawn_file_actions_init(&fa);
    posix_spawn_file_actions_addopen(&fa, 0, "/dev/null", O_RDONLY, 0);
    posix_spawn_file_actions_adddup2(&fa, pfd[1], 1);
    posix_spawn_file_actions_adddup2(&fa, pfd[1], 2);
    posix_spawn_file_actions_addclose(&fa, pfd[0]);
    posix_spawn_file_actions_addclose(&fa, pfd[1]);
    pid_t pid = -1;
    int rc = posix_spawnp(&pid, argv[0], &fa, NULL, argv, environ);
    posix_spawn_file_actions_destroy(&fa);
    close(pfd[1]);
    if (rc != 0) {
        close(pfd[0]);
        return -1;
    }
    size_t len = 0;
    char tmp[512];
    ssize_t r;
    while ((r = read(pfd[0], tmp, sizeof(tmp))) > 0) {
        if (out && len + 1 < out_cap) {
            size_t take = (size_t)r;
            if (len + take + 1 > out_cap) take = out_cap - 1 - len;
            memcpy(out + len, tmp, take);
            len += take;
        }
    }
    if (out && out_cap) out[len] = '\0';
    close(pfd[0]);
    int status = 0;
    while (waitpid(pid, &status, 0) < 0 && errno == EINTR) {
    }
    return WIFEXITED(status) ? WEXITSTATUS(status) : -1;
}

static int sh_c(const char* cmd, char* out, size_t out_cap) { /* exec.Command("sh", "-c", cmd).CombinedOutput() */
    char* argv[] = {(char*)"sh", (char*)"-c", (char*)cmd, NULL};
    return run_capture(argv, out, out_cap);
}

static int nsenter_sh_c(int target_pid, const char* cmd, char* out, size_t out_cap) {
    /* namespace.Config{Mount:true, Target:PID}.Execute("sh", "-c", cmd): util.go:60-64 + namespace.go:42-68 */
    gmo_nsenter_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.mount = 1;
    cfg.target = target_pid;
    char* argv[32];
    char pool[8192];
    const char* args[] = {"-c", cmd};
    int argc = gmo_nsenter_argv(&cfg, "sh", args, 2, argv, 32, pool, sizeof(pool));
    if (argc < 0) return argc;
    return run_capture(argv, out, out_cap);
}

static void strip_docker_prefix(const char* id, char* out, size_t cap) { /* strings.Replace(id, "docker://", "", 1) util.go:23 */
    const char* hit = strstr(id, "docker://");
    if (hit == NULL) {
        copy_str(out, cap, id);
        return;
    }
    sb s;
    sb_init(&s, out, cap);
    sb_putn(&s, id, (size_t)(hit - id));
    sb_put(&s, hit + strlen("docker://"));
}

static int go_atoi(const char* s, int* out) { /* strconv.Atoi: optional sign, digits only, no spaces */
    const char* p = s;
    if (*p == '+' || *p == '-') ++p;
    if (*p == '\0') return -1;
    for (const char* q = p; *q; ++q)
        if (!isdigit((unsigned char)*q)) return -1;
    *out = atoi(s);
    return 0;
}

static int resolve_cgroup(const gmo_mount_env* env, const gmo_pod* pod, char* cgroup_path, size_t cap, char* err, size_t err_cap) {
    char cid[512], driver[32];
    strip_docker_prefix(pod->container_id, cid, sizeof(cid));                            /* util.go:22-23 */
    int rc = gmo_get_cgroup_driver(env->cgroup_driver, driver, sizeof(driver), err, err_cap); /* util.go:25-29 */
    if (rc != 0) return rc;
    rc = gmo_get_cgroup_name(driver, pod, cid, cgroup_path, cap);                          /* util.go:30-34 */
    if (rc != 0 && err && err_cap) snprintf(err, err_cap, "Get cgroup path for Pod: %s failed", pod->name);
    return rc;
}

int gmo_mount_gpu(const gmo_mount_env* env, const gmo_pod* pod, const gmo_gpu* gpu, char* err, size_t err_cap) { /* util.go:17-71 */
    char cgroup_path[4096], cmd[8192], out[1024];
    if (err && err_cap) err[0] = '\0';
    int rc = resolve_cgroup(env, pod, cgroup_path, sizeof(cgroup_path), err, err_cap);
    if (rc != 0) return rc;
    /* cgroup.AddGPUDevicePermission, util.go:37 -> cgroup.go:143-155 */
    if (gmo_permission_command(env->devices_root, cgroup_path, gpu->minor_number, 1, cmd, sizeof(cmd)) != 0) return GMO_E_BUF;
    if (sh_c(cmd, out, sizeof(out)) != 0) {
        if (err && err_cap) snprintf(err, err_cap, "Exec \"%s\" failed: %s", cmd, out);
        return GMO_E_EXEC;
    }
    /* cgroup.GetCgroupPIDs, util.go:44-49 */
    char pids[1][16];
    int n = gmo_get_cgroup_pids(env->devices_root, cgroup_path, pids, 1);
    if (n < 0) {
        if (err && err_cap) snprintf(err, err_cap, "open cgroup.procs failed");
        return GMO_E_IO;
    }
    if (n == 0) return GMO_E_PANIC_INDEX; /* util.go:50 pids[0]: index out of range */
    int pid;
    if (go_atoi(pids[0], &pid) != 0) return GMO_E_ATOI; /* util.go:50-55 */
    /* namespace.AddGPUDeviceFile, util.go:60-67 -> namespace.go:167-177 */
    if (gmo_mknod_command(env->dev_prefix, gpu->minor_number, cmd, sizeof(cmd)) != 0) return GMO_E_BUF;
    rc = nsenter_sh_c(pid, cmd, out, sizeof(out));
    if (rc == GMO_E_TARGET) return rc;
    if (rc != 0) {
        if (err && err_cap) snprintf(err, err_cap, "Failed to execute cmd: %s: %s", cmd, out);
        return GMO_E_EXEC;
    }
    return GMO_E_OK;
}

int gmo_unmount_gpu(const gmo_mount_env* env, const gmo_pod* pod, const gmo_gpu* gpu, int force, const unsigned* gpu_pids,
                    int n_gpu_pids, char* err, size_t err_cap) { /* util.go:73-147 */
    char cgroup_path[4096], cmd[8192], out[1024];
    if (err && err_cap) err[0] = '\0';
    int rc = resolve_cgroup(env, pod, cgroup_path, sizeof(cgroup_path), err, err_cap);
    if (rc != 0) return rc;
    /* util.go:93-98 */
    enum { kMaxPids = 4096 };
    static __thread char pids[kMaxPids][16];
    int n = gmo_get_cgroup_pids(env->devices_root, cgroup_path, pids, kMaxPids);
    if (n < 0) return GMO_E_IO;
    if (n > kMaxPids) n = kMaxPids;
    /* GetPodGPUProcesses, util.go:100-105 -> :152-196 */
    const char* pid_ptrs[kMaxPids];
    for (int i = 0; i < n; ++i) pid_ptrs[i] = pids[i];
    int busy_idx[2048];
    if (n_gpu_pids > 2048) n_gpu_pids = 2048;
    int n_busy = gmo_pod_gpu_processes(pid_ptrs, n, gpu_pids, n_gpu_pids, busy_idx);
    if (n_busy != 0 && !force) return GMO_E_BUSY; /* util.go:106-109: errors.New(string(RemoveGPUResponse_GPUBusy)) */
    /* cgroup.RemoveGPUDevicePermission, util.go:112-115 -> cgroup.go:157-169 */
    if (gmo_permission_command(env->devices_root, cgroup_path, gpu->minor_number, 0, cmd, sizeof(cmd)) != 0) return GMO_E_BUF;
    if (sh_c(cmd, out, sizeof(out)) != 0) {
        if (err && err_cap) snprintf(err, err_cap, "Exec \"%s\" failed: %s", cmd, out);
        return GMO_E_EXEC;
    }
    if (n == 0) return GMO_E_PANIC_INDEX; /* util.go:118 pids[0] */
    int pid;
    if (go_atoi(pids[0], &pid) != 0) return GMO_E_ATOI;
    /* namespace.RemoveGPUDeviceFile, util.go:131-134 -> namespace.go:179-189 */
    if (gmo_rm_command(env->dev_prefix, gpu->minor_number, cmd, sizeof(cmd)) != 0) return GMO_E_BUF;
    rc = nsenter_sh_c(pid, cmd, out, sizeof(out));
    if (rc != 0) {
        if (err && err_cap) snprintf(err, err_cap, "Failed to execute cmd: %s: %s", cmd, out);
        return rc == GMO_E_TARGET ? rc : GMO_E_EXEC;
    }
    if (n_busy != 0) { /* util.go:137-142 -> namespace.go:191-201 */
        char pidstr[2048][16];
        const char* ptrs[2048];
        for (int i = 0; i < n_busy; ++i) {
            snprintf(pidstr[i], 16, "%d", (int)gpu_pids[busy_idx[i]]);
            ptrs[i] = pidstr[i];
        }
        char killcmd[40000];
        if (gmo_kill_command(ptrs, n_busy, killcmd, sizeof(killcmd)) != 0) return GMO_E_BUF;
        rc = nsenter_sh_c(pid, killcmd, out, sizeof(out));
        if (rc != 0) {
            if (err && err_cap) snprintf(err, err_cap, "Failed to execute cmd: %s: %s", killcmd, out);
            return GMO_E_EXEC;
        }
    }
    return GMO_E_OK;
}
