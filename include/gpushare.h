/*
 * gpushare.h -- flat C ABI of libgpushare_core.so
 *
 * B200-native core of the gpushare device plugin's per-node hot path:
 *   NVML enumerate -> fake-device expansion / ListAndWatch payload -> Allocate
 *   decision + env injection -> health re-advertise, plus the sm_100a HBM
 *   slice-probe kernel (new: the reference runs no GPU code).
 *
 * Every entry point names the reference interface it replaces as
 * `path:line` relative to the reference tree (cjg/aliyun-gpushare-device-plugin).
 * The reference binds NVML through go-nvml and speaks protobuf through gogo; a
 * Go maintainer would bind this header with cgo.
 *
 * Conventions
 *   - every function returns an int status: GS_OK (0) or a negative GS_E_* code;
 *     nothing throws or aborts across this boundary;
 *   - output buffers are caller-owned (pointer + capacity); when the capacity is
 *     too small the call returns GS_E_NOSPACE and still writes the required
 *     length to its *_len out-parameter;
 *   - strings are (pointer,length) pairs unless documented as NUL-terminated;
 *   - gs_table is immutable after gs_table_build except for its health bits,
 *     which are atomics: concurrent readers need no lock;
 *   - gs_allocate and gs_plugin_allocate perform no heap allocation in steady state
 *     (per-thread scratch grows once to the largest pod list seen);
 *   - a gs_probe serves one caller at a time (one batch in flight); use one per GPU;
 *   - pod *claiming* must be serialised by the caller exactly as the
 *     reference's m.Lock() does (allocate.go:60-61); gs_plugin_* does that
 *     internally with one mutex.
 */
#ifndef GPUSHARE_H
#define GPUSHARE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS_API __attribute__((visibility("default")))

/* ---- status codes ------------------------------------------------------- */
#define GS_OK 0
#define GS_E_INVAL (-1)     /* bad argument */
#define GS_E_NOSPACE (-2)   /* caller buffer too small; required size reported */
#define GS_E_NOTFOUND (-3)  /* unknown id / index */
#define GS_E_WIRE (-4)      /* malformed protobuf wire bytes */
#define GS_E_NVML (-5)      /* NVML call failed; see gs_nvml_last_error */
#define GS_E_CUDA (-6)      /* CUDA call failed; see gs_cuda_last_error */
#define GS_E_NOLIB (-7)     /* shared library could not be loaded */
#define GS_E_STATE (-8)     /* call made in the wrong state (e.g. not opened) */
#define GS_E_TIMEOUT (-9)   /* wait timed out (not an error for gs_health_wait) */
#define GS_E_LIMIT (-10)    /* size beyond a documented maximum */

GS_API const char *gs_strerror(int status);
GS_API const char *gs_version(void);

/* ---- constants of the drop-in surface (const.go:11-35) ------------------ */
#define GS_RESOURCE_NAME "aliyun.com/gpu-mem"                 /* const.go:11 */
#define GS_RESOURCE_COUNT "aliyun.com/gpu-count"              /* const.go:12 */
#define GS_SERVER_SOCK_NAME "aliyungpushare.sock"             /* const.go:13 */
#define GS_DEVICE_PLUGIN_PATH "/var/lib/kubelet/device-plugins/"
#define GS_KUBELET_SOCKET GS_DEVICE_PLUGIN_PATH "kubelet.sock"
#define GS_ENV_NVGPU "NVIDIA_VISIBLE_DEVICES"                 /* const.go:24 */
#define GS_ENV_RESOURCE_INDEX "ALIYUN_COM_GPU_MEM_IDX"        /* const.go:25 */
#define GS_ENV_RESOURCE_BY_POD "ALIYUN_COM_GPU_MEM_POD"       /* const.go:26 */
#define GS_ENV_RESOURCE_BY_CONTAINER "ALIYUN_COM_GPU_MEM_CONTAINER" /* :27 */
#define GS_ENV_RESOURCE_BY_DEV "ALIYUN_COM_GPU_MEM_DEV"       /* const.go:28 */
#define GS_ENV_ASSIGNED_FLAG "ALIYUN_COM_GPU_MEM_ASSIGNED"    /* const.go:29 */
#define GS_ENV_ASSUME_TIME "ALIYUN_COM_GPU_MEM_ASSUME_TIME"   /* const.go:30 */
#define GS_ENV_CGPU_DISABLE "CGPU_DISABLE"                    /* allocate.go:126 */
#define GS_NODE_LABEL_DISABLE_CGPU "cgpu.disable.isolation"   /* const.go:32 */
#define GS_HEALTHY "Healthy"      /* pluginapi.Healthy   (k8s.io/kubelet v0.28.3) */
#define GS_UNHEALTHY "Unhealthy"  /* pluginapi.Unhealthy */

/* MemoryUnit (const.go:8,34-35; cmd/nvidia/main.go:67-78) */
#define GS_UNIT_GIB 0
#define GS_UNIT_MIB 1
/* translatememoryUnits (cmd/nvidia/main.go:67-78): "MiB" -> GS_UNIT_MIB,
 * "GiB" -> GS_UNIT_GIB, anything else -> GS_UNIT_GIB with *warned = 1. */
GS_API int gs_translate_memory_unit(const char *s, size_t len, int *unit, int *warned);
GS_API const char *gs_unit_name(int unit); /* "GiB" / "MiB" */

/* ---- NVML / driver layer ------------------------------------------------- *
 * Replaces the go-nvml v0.12.0-1 call sites: nvml.Init/Shutdown
 * (gpumanager.go:36,41), DeviceGetCount (nvidia.go:45,51), DeviceGetHandleByIndex
 * (:57), GetUUID (:60), GetMemoryInfo_v2 (:64), ErrorString (:19). The library
 * is dlopen'ed at run time so the same .so loads on GPU-less machines and
 * against the fake NVML used for BASELINE config 1/5.                        */
#define GS_UUID_MAX 96

typedef struct gs_gpu_info {
  char uuid[GS_UUID_MAX];   /* NUL-terminated, e.g. "GPU-xxxxxxxx-xxxx-..." */
  char pci_bus_id[32];      /* NUL-terminated "00000000:1B:00.0" or ""      */
  uint64_t total_bytes;     /* nvmlMemory_v2_t.total (bytes, nvml.h:257-264) */
  int32_t nvml_index;
  int32_t cuda_ordinal;     /* -1 until gs_cuda_map_ordinals resolves it     */
} gs_gpu_info;

/* lib_path NULL -> "libnvidia-ml.so.1". Calls nvmlInit_v2. */
GS_API int gs_nvml_open(const char *lib_path);
GS_API int gs_nvml_close(void); /* nvmlShutdown + dlclose */
GS_API int gs_nvml_device_count(uint32_t *count);          /* nvidia.go:44-48 */
GS_API int gs_nvml_device_info(uint32_t index, gs_gpu_info *out); /* :57-64   */
GS_API const char *gs_nvml_last_error(void); /* nvmlErrorString of last failure */
/* Fills cuda_ordinal of each entry by PCI bus id (NVML index != CUDA ordinal
 * in general). Needs a CUDA driver; returns GS_E_CUDA without one.           */
GS_API int gs_cuda_map_ordinals(gs_gpu_info *gpus, uint32_t n);
GS_API const char *gs_cuda_last_error(void);

/* Health watch: the rule of the commented-out watchXIDs (nvidia.go:97-153).
 * gs_health_open registers nvmlEventTypeXidCriticalError on every GPU; GPUs
 * that answer NOT_SUPPORTED are reported in *unsupported_mask (the reference
 * marks those unhealthy at once, nvidia.go:107-111).                          */
#define GS_ALL_GPUS (-1)
typedef struct gs_health_event {
  int32_t gpu_index; /* NVML index, or GS_ALL_GPUS when the event has no UUID */
  int32_t ignored;   /* 1 for XID 31/43/45 (application errors, nvidia.go:134) */
  uint64_t xid;
} gs_health_event;
GS_API int gs_health_open(uint64_t *unsupported_mask);
/* GS_OK with *out filled, or GS_E_TIMEOUT after timeout_ms (reference: 5000). */
GS_API int gs_health_wait(uint32_t timeout_ms, gs_health_event *out);
GS_API int gs_health_close(void);

/* ---- device model: fake-ID expansion (nvidia.go:23-95) ------------------- */
typedef struct gs_table gs_table;

/* setGPUMemory (nvidia.go:31-38): units = raw (MiB unit) or raw/1024 (GiB).
 * `raw` is the MiB-valued total of GPU 0 (total_bytes >> 20).      */
GS_API uint64_t gs_units_from_raw(uint64_t raw_mib, int unit);

/* getDevices (nvidia.go:50-86): for NVML index i = 0..n-1, `units` fake
 * devices "<uuid>-_-<j>", j = 0..units-1, (gpu asc, j asc). Every GPU gets
 * the unit count derived from GPU 0 (H7). uuids[i] need not be NUL-terminated. */
GS_API int gs_table_build(const char *const *uuids, const uint32_t *uuid_lens,
                          uint32_t n_gpus, uint64_t units_per_gpu, int unit,
                          gs_table **out);
GS_API int gs_table_build_from_gpus(const gs_gpu_info *gpus, uint32_t n_gpus,
                                    int unit, gs_table **out);
GS_API void gs_table_free(gs_table *t);
GS_API uint32_t gs_table_gpu_count(const gs_table *t);     /* len(devNameMap) */
GS_API uint64_t gs_table_units_per_gpu(const gs_table *t); /* getGPUMemory()  */
GS_API uint64_t gs_table_device_count(const gs_table *t);  /* len(m.devs)     */
GS_API int gs_table_unit(const gs_table *t);

/* generateFakeDeviceID (nvidia.go:23-25). Returns length via *len. */
GS_API int gs_fake_id_format(const gs_table *t, uint32_t gpu, uint64_t unit_index,
                             char *buf, size_t cap, size_t *len);
/* extractRealDeviceID (nvidia.go:27-29): length of the text before the first
 * "-_-" (whole string when absent, as strings.Split(...)[0]).                 */
GS_API size_t gs_extract_real_device_id(const char *id, size_t len);
/* O(1) fake-ID -> (gpu, unit): hashes the UUID prefix, parses the decimal
 * suffix. GS_E_NOTFOUND for ids this table never advertised.                  */
GS_API int gs_id_lookup(const gs_table *t, const char *id, size_t len,
                        uint32_t *gpu, uint64_t *unit_index);
/* GetDeviceNameByIndex (server.go:80-91). Borrowed pointer, valid for t's life. */
GS_API int gs_device_name_by_index(const gs_table *t, uint64_t index,
                                   const char **uuid, size_t *len);
/* devNameMap lookup (server.go:22): uuid -> NVML index. */
GS_API int gs_device_index_by_name(const gs_table *t, const char *uuid, size_t len,
                                   uint32_t *index);

/* ---- ListAndWatch (server.go:180-197) ------------------------------------ *
 * Health is per fake device as in the reference (d.Health = Unhealthy,
 * server.go:189) with a per-GPU fast path; there is no recovery in the
 * reference (FIXME server.go:188) -- gs_law_reset_health exists for restart.   */
GS_API int gs_law_mark_unhealthy_device(gs_table *t, uint32_t gpu, uint64_t unit_index);
GS_API int gs_law_mark_unhealthy_id(gs_table *t, const char *id, size_t len);
GS_API int gs_law_mark_unhealthy_gpu(gs_table *t, int32_t gpu /* or GS_ALL_GPUS */);
GS_API int gs_law_reset_health(gs_table *t);
GS_API int gs_law_device_healthy(const gs_table *t, uint32_t gpu, uint64_t unit_index);
GS_API uint64_t gs_law_generation(const gs_table *t); /* bumps on every change */

/* Exact size in bytes of the v1beta1.ListAndWatchResponse for the current
 * health state: sum over devices of 13 + len(id) (+2 per unhealthy device).   */
GS_API size_t gs_law_payload_size(const gs_table *t);
/* Encodes the whole ListAndWatchResponse{devices=1} into buf.                */
GS_API int gs_law_encode(const gs_table *t, uint8_t *buf, size_t cap, size_t *len);
/* Zero-copy form: borrowed pre-encoded segments (one per GPU) whose concatenation is the
 * response. A segment of a GPU with unhealthy devices is rebuilt by the next gs_law_payload
 * call after a health change, so this form is for ONE sender (kubelet opens one stream);
 * several concurrent senders use gs_law_encode, which copies under the table's lock.
 * Segments of all-healthy GPUs never move until gs_table_free.                           */
typedef struct gs_iovec {
  const uint8_t *base;
  size_t len;
} gs_iovec;
GS_API int gs_law_payload(gs_table *t, gs_iovec *iov, uint32_t cap, uint32_t *n);

/* ---- Allocate (allocate.go:25-201, podmanager.go:247-347, podutils.go) ---- */
/* One pod as the caller's lister returned it (kubelet /pods or apiserver
 * LIST). Absent annotation <=> pointer NULL (length ignored). Strings are not
 * NUL-terminated.                                                             */
typedef struct gs_str {
  const char *p;
  size_t n;
} gs_str;

typedef struct gs_pod_view {
  gs_str uid;        /* metadata.uid        (dedup key, podmanager.go:287-290) */
  gs_str ns;         /* metadata.namespace                                    */
  gs_str name;       /* metadata.name                                         */
  gs_str node_name;  /* spec.nodeName       (podmanager.go:274)               */
  /* Sum over spec.containers of uint(Limits["aliyun.com/gpu-mem"].Value())
   * with Go's wrapping uint add (podutils.go:122-131); see gs_quantity_value. */
  uint64_t gpu_mem_limit;
  /* podutils.go:41 guards on len(annotations) > 0; a pod that reaches that line
   * was selected by its annotations, so the three values below carry it all.   */
  gs_str ann_idx;          /* ALIYUN_COM_GPU_MEM_IDX          (NULL = absent) */
  gs_str ann_assume_time;  /* ALIYUN_COM_GPU_MEM_ASSUME_TIME  (NULL = absent) */
  gs_str ann_assigned;     /* ALIYUN_COM_GPU_MEM_ASSIGNED     (NULL = absent) */
} gs_pod_view;

typedef struct gs_alloc_ctx {
  const gs_table *table;
  gs_str node_name;              /* NODE_NAME (podmanager.go:55)               */
  int32_t disable_cgpu_isolation;/* node label cgpu.disable.isolation == "true" */
  int64_t now_unix_nano;         /* injected clock for the patch body (H4)     */
} gs_alloc_ctx;

/* decision kinds */
#define GS_ALLOC_MATCHED 0        /* allocate.go:91-152: envs + patch          */
#define GS_ALLOC_SINGLE_GPU 1     /* allocate.go:154-180: UUID fallback        */
#define GS_ALLOC_ERR_NO_MATCH 2   /* allocate.go:182-187                       */
#define GS_ALLOC_ERR_BAD_INDEX 3  /* allocate.go:93-111 (id < 0 or unknown)    */
#define GS_ALLOC_ERR_PODS 4       /* allocate.go:64-67 (lister failed)         */
#define GS_ALLOC_ERR_PATCH 5      /* allocate.go:133-151 (patch failed)        */

typedef struct gs_alloc_result {
  int32_t kind;
  int32_t pod_index;     /* index into pods[] of the assumed pod, else -1     */
  int32_t gpu_index;     /* NVML index granted, else -1                        */
  uint32_t n_containers; /* len(reqs.ContainerRequests)                        */
  uint64_t pod_req_units;/* podReqGPU (allocate.go:55-57)                      */
  size_t resp_len;       /* bytes of v1beta1.AllocateResponse written          */
  size_t patch_len;      /* bytes of the strategic-merge patch JSON (0 if none) */
} gs_alloc_result;

/* The whole decision for one AllocateRequest (wire bytes as kubelet sent
 * them): count ids per container without materialising them, filter and
 * order the candidate pods, match, read the GPU index, encode the response
 * and the patch body (podutils.go:27-35, keys in json.Marshal order).
 * Never fails for a "bad" pod: like the reference every such outcome is a
 * normal response carrying the poison envs of buildErrResponse (H8).          */
GS_API int gs_allocate(const gs_alloc_ctx *ctx, const uint8_t *req, size_t req_len,
                       const gs_pod_view *pods, size_t n_pods,
                       uint8_t *resp, size_t resp_cap,
                       char *patch, size_t patch_cap, gs_alloc_result *out);
/* buildErrResponse (allocate.go:25-40) for callers whose lister or PATCH
 * failed (allocate.go:64-67,133-151).                                          */
GS_API int gs_allocate_error_response(const gs_alloc_ctx *ctx, const uint8_t *req,
                                      size_t req_len, int32_t kind, uint8_t *resp,
                                      size_t resp_cap, gs_alloc_result *out);
/* Per-container id counts of an AllocateRequest (allocate.go:55-57,115). */
GS_API int gs_request_scan(const uint8_t *req, size_t req_len, uint32_t *counts,
                           uint32_t cap, uint32_t *n_containers, uint64_t *total);

/* Go-semantics helpers the lister side needs to fill gs_pod_view. */
/* resource.Quantity parse + Value() (apimachinery v0.28.4): integer, rounded
 * up; suffixes m,k,M,G,T,P,E,Ki,Mi,Gi,Ti,Pi,Ei and decimal exponents.         */
GS_API int gs_quantity_value(const char *s, size_t len, int64_t *value);
GS_API int gs_go_atoi(const char *s, size_t len, int64_t *value);        /* strconv.Atoi */
GS_API int gs_go_parse_uint64(const char *s, size_t len, uint64_t *value); /* ParseUint(s,10,64) */

/* ---- host-side plugin mirror (server.go:19-35 NvidiaDevicePlugin) --------- *
 * Owns the lock the reference takes in Allocate, an in-memory pod source that
 * stands in for the lister/patcher (network I/O stays with the caller), and
 * optionally a probe engine per GPU.                                           */
typedef struct gs_plugin gs_plugin;
GS_API int gs_plugin_new(gs_table *table, const char *node_name, size_t node_len,
                         int32_t disable_cgpu_isolation, gs_plugin **out);
GS_API void gs_plugin_free(gs_plugin *p);
/* Replace the pod source contents (deep copy). */
GS_API int gs_plugin_set_pods(gs_plugin *p, const gs_pod_view *pods, size_t n);
GS_API int gs_plugin_pod_count(const gs_plugin *p, size_t *n);
/* Read back one pod's annotations after patches were applied. */
GS_API int gs_plugin_get_pod(const gs_plugin *p, size_t index, gs_pod_view *out);
GS_API int gs_plugin_set_clock(gs_plugin *p, int64_t fixed_unix_nano /* 0 = real */);
/* Allocate (allocate.go:43-201) end to end against the in-memory pod source:
 * decision, response bytes, and the patch applied to the source.              */
GS_API int gs_plugin_allocate(gs_plugin *p, const uint8_t *req, size_t req_len,
                              uint8_t *resp, size_t resp_cap, gs_alloc_result *out);
/* GetDevicePluginOptions / PreStartContainer (server.go:93-95,199-201): empty
 * messages; resp_len is always 0.                                              */
GS_API int gs_plugin_get_device_plugin_options(gs_plugin *p, uint8_t *resp,
                                               size_t cap, size_t *len);
GS_API int gs_plugin_pre_start_container(gs_plugin *p, const uint8_t *req,
                                         size_t req_len, uint8_t *resp, size_t cap,
                                         size_t *len);
/* ListAndWatch pump (server.go:180-193): blocks until the health generation
 * differs from *seen_generation or stop is requested; returns GS_E_TIMEOUT
 * after timeout_ms. On GS_OK the caller re-sends gs_law_payload.               */
GS_API int gs_plugin_wait_health_change(gs_plugin *p, uint64_t *seen_generation,
                                        uint32_t timeout_ms);
GS_API int gs_plugin_notify_health(gs_plugin *p); /* wake waiters after gs_law_mark_* */
GS_API int gs_plugin_stop(gs_plugin *p);          /* close(m.stop), server.go:152 */
GS_API int gs_plugin_stopped(const gs_plugin *p);

/* ---- kubelet-facing transport (server.go:114-177,232-265) ------------------ *
 * gRPC over HTTP/2 on a unix socket, native: the reference gets it from
 * google.golang.org/grpc (go.mod:48). The server routes v1beta1.DevicePlugin
 * to the gs_plugin_* entry points above on one serving thread; ListAndWatch
 * sends borrow gs_law_payload's segments, so while a server runs on a table it
 * is that table's one gs_law_payload caller (others use gs_law_encode).
 * Limits: 64 connections, 128 open streams on each (advertised and enforced);
 * request messages up to max_request_bytes (default 64 MiB; grpc-go's
 * default of 4 MiB would refuse a whole B200 at MiB granularity, H12), header
 * blocks up to 256 KiB, Huffman-coded header strings restricted to printable
 * ASCII (see csrc/hpack.h), no message compression (UNIMPLEMENTED, as grpc-go
 * answers without a registered decompressor).                                  */
typedef struct gs_server gs_server;
/* Optional replacement for the Allocate body, for callers whose pod source is
 * the network (podmanager.go:187-245, allocate.go:136): called with the
 * AllocateRequest bytes; writes the AllocateResponse bytes. Calls are made one
 * at a time, in arrival order, on a worker thread of the server (the callback
 * may block on the network for seconds; ListAndWatch and other connections
 * keep being served meanwhile). GS_E_NOSPACE with *resp_len = required size
 * asks for a larger buffer (the callback is then called again: keep the answer). */
typedef int (*gs_server_allocate_fn)(void *user, const uint8_t *req, size_t req_len,
                                     uint8_t *resp, size_t resp_cap, size_t *resp_len);
/* Optional hook after each decision of the built-in Allocate (e.g. queue a
 * slice probe for the granted range). Runs on the serving thread.              */
typedef void (*gs_server_decision_fn)(void *user, const gs_alloc_result *decision);
typedef struct gs_server_config {
  const char *socket_path;        /* e.g. GS_DEVICE_PLUGIN_PATH GS_SERVER_SOCK_NAME */
  gs_server_allocate_fn allocate; /* NULL: gs_plugin_allocate on the in-memory pod store */
  gs_server_decision_fn on_decision;
  void *user;
  size_t max_request_bytes;       /* 0 = 64 MiB */
} gs_server_config;
typedef struct gs_server_stats {
  uint64_t connections, requests, allocate_calls, law_messages, protocol_errors;
  uint64_t bytes_in, bytes_out;   /* of connections already closed */
  uint64_t recv_body_ns;          /* spent in recv() of request bodies (DATA frames)          */
  uint64_t allocate_ns;           /* spent in the built-in Allocate: scan + decision + encode  */
} gs_server_stats;
/* Start (server.go:114-142): removes a stale socket, listens, serves, and dials
 * itself (5 s) before returning. GS_E_STATE: cannot bind; GS_E_TIMEOUT: the
 * self-dial failed.                                                            */
GS_API int gs_server_start(gs_plugin *p, const gs_server_config *cfg, gs_server **out);
/* Stop (server.go:145-155): gs_plugin_stop (open ListAndWatch streams end with
 * status OK), GOAWAY, close, remove the socket, free the server.               */
GS_API int gs_server_stop(gs_server *s);
GS_API int gs_server_get_stats(const gs_server *s, gs_server_stats *out);

/* Register (server.go:158-177): dial kubelet's socket (blocking, timeout), send
 * RegisterRequest{version "v1beta1", endpoint, resource_name}, options unset.
 * GS_E_TIMEOUT: no kubelet; GS_E_STATE: non-OK status (gs_grpc_last_error).    */
GS_API int gs_register(const char *kubelet_socket, const char *endpoint,
                       const char *resource_name, uint32_t timeout_ms);

/* The kubelet side of the same transport, for Register above, tests and
 * benchmarks: unary calls and server streams over one connection. One caller at
 * a time per client. No counterpart in the reference beyond dial()
 * (server.go:252-265).                                                         */
typedef struct gs_grpc_client gs_grpc_client;
GS_API int gs_grpc_client_connect(const char *unix_path, uint32_t timeout_ms,
                                  gs_grpc_client **out);
GS_API void gs_grpc_client_close(gs_grpc_client *c);
/* GS_OK when the call completed; *grpc_status then carries the gRPC code (0 =
 * OK, response in resp) and `message` the grpc-message, NUL-terminated.        */
GS_API int gs_grpc_client_unary(gs_grpc_client *c, const char *method_path,
                                const uint8_t *req, size_t req_len, uint8_t *resp,
                                size_t resp_cap, size_t *resp_len, int32_t *grpc_status,
                                char *message, size_t message_cap);
GS_API int gs_grpc_client_stream_open(gs_grpc_client *c, const char *method_path,
                                      const uint8_t *req, size_t req_len, uint32_t *stream);
/* Next message of a server stream. GS_OK: a message; GS_E_NOSPACE: it needs
 * *resp_len bytes (still queued); GS_E_TIMEOUT; GS_E_STATE: the stream ended,
 * *grpc_status holds its status (or -1 when the connection broke).             */
GS_API int gs_grpc_client_stream_next(gs_grpc_client *c, uint32_t stream, uint8_t *resp,
                                      size_t resp_cap, size_t *resp_len,
                                      int32_t *grpc_status, uint32_t timeout_ms);
/* Same, without the copy: *msg points at the bytes where they were received and
 * stays valid until the next call on this client.                             */
GS_API int gs_grpc_client_stream_next_borrow(gs_grpc_client *c, uint32_t stream,
                                             const uint8_t **msg, size_t *len,
                                             int32_t *grpc_status, uint32_t timeout_ms);
GS_API int gs_grpc_client_stream_cancel(gs_grpc_client *c, uint32_t stream);
GS_API const char *gs_grpc_last_error(void); /* thread-local, valid until the next call */

/* ---- slice probe (K1; no counterpart in the reference) -------------------- *
 * Writes pattern(i, seed) over exactly the granted HBM range as 128-bit
 * coalesced stores, reads it back with 128-bit loads, compares, and reduces
 * errors with a warp ballot. Two passes => 2 algorithmic bytes per slice byte. */
typedef struct gs_probe gs_probe;

/* pattern(i, seed) written to and expected in the slice's i-th 16-byte word.
 *   LOWTOGGLE  { r0 ^ lo(i), r1 ^ hi(i), r2, r3 }, r0..r3 drawn from the seed once per sweep:
 *              every cell gets a seed-dependent bit and every word its own address, while
 *              few bits change from word to word -- which is what the power-capped store
 *              pass is priced by. The fast form; what grants are swept with.
 *   HASH       { h, ~h, h * K + lo(i), rotl(h, 13) ^ hi(i) ^ C }, h an avalanche hash of
 *              (i, seed): every lane carries the address, both polarities in one sweep. The
 *              strong form; what health-time sweeps use. Costs ~9 % of the store rate.  */
#define GS_PROBE_PATTERN_LOWTOGGLE 0
#define GS_PROBE_PATTERN_HASH 1

typedef struct gs_slice {
  void *base;        /* device pointer, 16-byte aligned                        */
  uint64_t bytes;    /* multiple of 16                                         */
  uint32_t seed;     /* change it every sweep: a stale sweep then never matches */
  uint32_t tag;      /* caller cookie (e.g. pod index); not interpreted        */
  uint32_t pattern;  /* GS_PROBE_PATTERN_*; anything else => GS_E_INVAL        */
  uint32_t reserved; /* 0                                                      */
} gs_slice;

typedef struct gs_slice_result {
  uint64_t mismatches;     /* 16-byte words that differ                        */
  uint64_t first_bad_byte; /* lowest differing byte offset; UINT64_MAX if none  */
  uint64_t bit_flips;      /* popcount(got ^ expected) over the slice           */
  uint64_t reserved;
} gs_slice_result;

typedef struct gs_probe_timing {
  float write_ms;    /* CUDA-event time of the store pass (0 if not run)       */
  float verify_ms;   /* CUDA-event time of the load+compare pass               */
  uint64_t bytes;    /* sum of slice bytes in the batch                        */
  uint32_t launches; /* kernels launched for the batch                         */
  float scrub_ms;    /* CUDA-event time of the L2 scrub between the passes     */
} gs_probe_timing;

#define GS_PROBE_MAX_SLICES 64 /* per batch (descriptors travel as kernel params) */

GS_API int gs_probe_open(int cuda_ordinal, gs_probe **out);
GS_API int gs_probe_close(gs_probe *p);
GS_API int gs_probe_sm_count(const gs_probe *p, int *sms);
/* Granted-range memory: plain cudaMalloc/cudaFree on the probe's device. */
GS_API int gs_probe_mem_alloc(gs_probe *p, uint64_t bytes, void **dptr);
GS_API int gs_probe_mem_free(gs_probe *p, void *dptr);
GS_API int gs_probe_mem_info(gs_probe *p, uint64_t *free_bytes, uint64_t *total_bytes);
/* A seed no earlier sweep of this engine used: a nonce drawn when the engine was opened
 * plus an odd multiple of a per-engine counter (2^32 sweeps before the first repeat). The
 * pattern's only defence against a store pass that silently did nothing is that the data
 * left by an earlier sweep never matches -- which holds only if every sweep gets a seed of
 * its own; take it from here rather than from anything that can repeat (a pod index, a
 * constant). One caller at a time, like the rest of a gs_probe.                        */
GS_API int gs_probe_next_seed(gs_probe *p, uint32_t *seed);
/* Asynchronous on the probe's stream. passes: bit0 = write, bit1 = verify, bit2 = scrub. */
#define GS_PROBE_WRITE 1u
#define GS_PROBE_VERIFY 2u
/* Between the passes, stream 192 MiB through the 126 MB L2 so the read-back comes from
 * DRAM even for slices smaller than the L2 (costs ~30 us and 192 MiB of HBM, allocated on
 * first use). Meaningful together with at least one of the two passes.               */
#define GS_PROBE_SCRUB 4u
GS_API int gs_probe_submit(gs_probe *p, const gs_slice *slices, uint32_t n,
                           uint32_t passes);
/* Waits for the last submit; copies the per-slice results (n as submitted). */
GS_API int gs_probe_wait(gs_probe *p, gs_slice_result *results, uint32_t n,
                         gs_probe_timing *timing);
GS_API int gs_probe_poll(gs_probe *p); /* GS_OK when idle, GS_E_TIMEOUT when busy */
/* submit + wait. */
GS_API int gs_probe_run(gs_probe *p, const gs_slice *slices, uint32_t n,
                        uint32_t passes, gs_slice_result *results,
                        gs_probe_timing *timing);
/* SURVEY 8(b) gs_probe_run: allocate `bytes` on the GPU, probe, free. */
GS_API int gs_probe_run_granted(int cuda_ordinal, uint64_t bytes, uint32_t seed,
                                gs_slice_result *result, gs_probe_timing *timing);
/* Test hook: XOR `mask` into the 32-bit word at byte_offset of a slice. */
GS_API int gs_probe_inject(gs_probe *p, void *base, uint64_t byte_offset,
                           uint32_t mask);
/* Test hook: while on, the store pass of this engine launches nothing, as a store kernel that
 * silently failed would. An engine opened with GS_PROBE_TEST_SKIP_STORES=n in the environment
 * runs its first n-1 store passes and drops the rest (how the daemon's sweeps are put to the
 * same test).                                                                               */
GS_API int gs_probe_test_skip_stores(gs_probe *p, int on);
/* Copy helpers so tests need no other CUDA binding. */
GS_API int gs_probe_read(gs_probe *p, const void *dptr, void *host, uint64_t bytes);
/* Kernel variant selection for tuning; -1 keeps the default.
 *   write_variant  100*T + 10*S + B : bulk async stores (default 102). T = 1/2/3 -> 256/512/1024
 *                  threads; S = 0/1/2 -> 3-stage ring / 2-stage / 3-stage + L2 evict-first hint;
 *                  B = 0..4 -> 4/2/8/16/1 KiB per bulk op
 *                  400 + B : paced plain STG.128, B = 0..6 -> 4/2/8/16/1/32/64 KiB per token
 *                  10*U + F : unpaced STG.128, U = 0/1/2 -> unroll 4/2/8, F = 0..3 -> plain/.cs/
 *                  L1::no_allocate/L2 evict-first
 *   verify_variant 10*U + F (default 10): U as above; F = 0..4 -> plain/nc+L1::no_allocate/.cs/
 *                  nc+L2 evict-first/nc+L2::256B prefetch
 *                  500 + 100*H + 10*C + W : bulk async loads into a shared-memory ring, compare
 *                  from shared memory. H = 0/1 -> plain / L2 evict-first hint; C = 0/1/2 ->
 *                  16/32/64 KiB stages; W = 0/1/2 -> 4/8/16 consumer warps
 *   ctas_per_sm    grid size of the STG/LDG kernels in CTAs per SM; 0 = resident CTAs (default).
 *                  For the bulk-load verify forms: 2 = two CTAs per SM with half the ring each,
 *                  anything else = one                                                          */
GS_API int gs_probe_set_variant(gs_probe *p, int write_variant, int verify_variant,
                                int ctas_per_sm);
/* Offered rate (GB/s, whole GPU) the store pass paces itself to; 0 = unpaced.
 * B200's write path sustains more when not over-offered. Default: 7600, fixed.
 * gs_probe_set_write_rate_auto(p, 1) lets the target back off for a board that
 * cannot keep up (down 0.8 % after a store pass of >= 4 GiB that fell more than
 * 1 % behind, back up 0.05 % after one that kept up, never above 7600); off by
 * default: on healthy boards the fixed target measured better (csrc/probe.cu).
 * Setting a rate here fixes it and turns the following off.                     */
GS_API int gs_probe_set_write_rate(gs_probe *p, uint32_t gbps);
GS_API int gs_probe_set_write_rate_auto(gs_probe *p, int on);
GS_API int gs_probe_get_write_rate(const gs_probe *p, uint32_t *gbps, int *is_auto);
/* Read-back pass for callers that may not allocate. No counterpart in the reference
 * (it runs no GPU code). The plugin's health checker on a GPU whose memory is granted
 * to tenants to the last unit cannot cudaMalloc (it would fail, or take what a tenant
 * was promised), and some callers cannot afford cudaMalloc/cudaFree's device-wide
 * synchronisation; this call therefore allocates nothing: the per-slice verdict table
 * is a fixed 2 KiB array resident on every GPU that loads the library, which is why
 * at most GS_PROBE_NOALLOC_MAX_SLICES slices fit (more => GS_E_LIMIT).
 *   - same load+compare kernel and verdicts as gs_probe_run(..., GS_PROBE_VERIFY, ...);
 *     the ranges must have been written before (gs_probe_submit with GS_PROBE_WRITE);
 *   - runs on `stream` (a cudaStream_t; NULL = the default stream) of device
 *     `cuda_ordinal` and returns after that stream has drained; no gs_probe needed;
 *     the verdicts come back through a cudaMemcpyAsync into ordinary (pageable) host
 *     memory, which the runtime stages and completes before returning -- the call
 *     is synchronous by design and pins nothing;
 *   - the first call on each device looks up the array's device address; make that
 *     first call before starting any stream capture;
 *   - every call on a device uses the same array and clears the rows it needs; two
 *     calls that overlap on one GPU race on it -- serialise them.                    */
#define GS_PROBE_NOALLOC_MAX_SLICES 64
GS_API int gs_probe_verify_noalloc(int cuda_ordinal, void *stream, const gs_slice *slices,
                                   uint32_t n, gs_slice_result *results);
/* Host restatement of the pattern is NOT here: it lives in oracle/pattern.c. */

#ifdef __cplusplus
}
#endif
#endif /* GPUSHARE_H */
