/*
 * igpu_rollup.h — C ABI of libigpu_rollup.so: the B200 (sm_100a) fleet rollup behind the
 * exported helpers of the reference's src/api/k8s.ts and src/api/metrics.ts.
 *
 * The reference (a TypeScript Headlamp plugin) has no native boundary today; these entry points
 * are what an N-API addon / ctypes stub for its rollup path binds (INTEGRATION.md shows both).
 * Every declaration cites the reference code whose arithmetic it replaces.  Citations are
 * relative to the reference repository root.
 *
 * Conventions
 *  - Plain C types only.  Every function returns an int32 status (IGPU_OK == 0, errors < 0) and
 *    never throws; igpu_last_error(ctx) has the text.  This mirrors the reference's "never throw
 *    on bad data" convention (k8s.ts:83,126,251; metrics.ts:73-74,98).
 *  - The caller owns every host buffer for the duration of a call; the library owns device
 *    memory behind the opaque handles.  No host pointer is retained after a call returns.
 *  - A context is bound to one CUDA device and one stream and is NOT thread-safe.
 *  - There is no CPU fallback: without a usable CUDA device igpu_ctx_create fails with
 *    IGPU_ERR_NODEVICE (north star: "the CUDA path has no CPU fallback").
 *  - Strings never cross this boundary: node names, resource keys, label keys, chip and instance
 *    labels are dictionary-encoded to integer ids by the host marshal layer, which also applies
 *    parseInt / parseFloat (k8s.ts:176,296; metrics.ts:130,137).
 *  - Integer results are exact (int64 on the device equals the reference's float64 sums while
 *    |sum| < 2^53; the marshal layer enforces that domain).  Float statistics are within 1e-5
 *    relative of the float64 definition stated per function.
 */
#ifndef IGPU_ROLLUP_H
#define IGPU_ROLLUP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IGPU_VERSION_MAJOR 0
#define IGPU_VERSION_MINOR 6 /* 0.2 series rollup, 0.3 native marshal + static join, 0.4 counters/rate + matrix ingest,
                                0.5 container ordinal in igpu_entry, per-container key rows, fleet quantile,
                                0.6 gap statistic of a counter slab, kept window counts (window rollup without a read) */
#define IGPU_VERSION_PATCH 0

/* ---- status codes ------------------------------------------------------------------------ */
#define IGPU_OK 0
#define IGPU_ERR_ARG (-1)      /* null pointer, inconsistent sizes, unsorted offsets ...        */
#define IGPU_ERR_CUDA (-2)     /* a CUDA runtime call failed; see igpu_last_error               */
#define IGPU_ERR_NCCL (-3)     /* NCCL missing or a collective failed                           */
#define IGPU_ERR_OOM (-4)      /* device or pinned-host allocation failed                       */
#define IGPU_ERR_NODEVICE (-5) /* no CUDA device / wrong architecture                           */
#define IGPU_ERR_RANGE (-6)    /* a size or id is beyond a documented limit                     */
#define IGPU_ERR_STATE (-7)    /* call order violated (e.g. download before rollup)             */

typedef struct igpu_ctx igpu_ctx;
typedef struct igpu_fleet igpu_fleet;
typedef struct igpu_series igpu_series;

/* ---- columnar encodings (host marshal layer <-> device) ------------------------------------ */

/* node_flags bits.  b0..b2: label === 'true' tests of k8s.ts:136-138 / :187-190.
 * b3: some capacity key starts with 'gpu.intel.com/' (k8s.ts:146-148).  b4: isNodeReady
 * (k8s.ts:329-331).  b5/b6: parseInt() without `|| 0` produced NaN while summing capacity /
 * allocatable (OverviewPage.tsx:98-99, NodesPage.tsx:173-174).                               */
#define IGPU_NODE_DISCRETE_ROLE 0x01u
#define IGPU_NODE_IGPU_ROLE 0x02u
#define IGPU_NODE_NFD_LABEL 0x04u
#define IGPU_NODE_HAS_GPU_CAPACITY 0x08u
#define IGPU_NODE_READY 0x10u
#define IGPU_NODE_CAP_NAN 0x20u
#define IGPU_NODE_ALLOC_NAN 0x40u

/* pod_phase codes (OverviewPage.tsx:122-130).  PROTO = a phase string that names an inherited
 * Object.prototype property ("toString", "constructor", ...): the reference's `phase in counts`
 * test is true for those, so such a pod lands in none of the five buckets.                    */
#define IGPU_PHASE_RUNNING 0
#define IGPU_PHASE_PENDING 1
#define IGPU_PHASE_SUCCEEDED 2
#define IGPU_PHASE_FAILED 3
#define IGPU_PHASE_OTHER 4
#define IGPU_PHASE_PROTO 5

#define IGPU_NODE_NONE 0xFFFFFFFFu /* pod without a (known) spec.nodeName                      */

/* Resource dictionary: ids 0..3 are fixed (k8s.ts:17-20); the marshal layer hands out ids >= 4
 * to any other 'gpu.intel.com/' key it meets ('tiles', custom names).                        */
#define IGPU_RES_I915 0
#define IGPU_RES_XE 1
#define IGPU_RES_MILLICORES 2
#define IGPU_RES_MEMORY_MAX 3
#define IGPU_MAX_RESOURCES 32

/* One entry per 'gpu.intel.com/' key of one container's requests or limits map.
 * value = parseInt(v, 10) || 0 (k8s.ts:296).  Entries are grouped by pod, pods ascending; inside a
 * pod they are grouped by container — spec.containers in order, then spec.initContainers
 * (k8s.ts:255) — and inside a container the requests map comes before the limits map, each in
 * the object's own key order (the order `new Set([...])` iterates, PodsPage.tsx:59-62).          */
#define IGPU_ENTRY_LIMIT 0x01u  /* key came from `limits` (else `requests`)                    */
#define IGPU_ENTRY_INIT 0x02u   /* container is an initContainer (k8s.ts:255,291)              */
#define IGPU_ENTRY_TRUTHY 0x04u /* the raw value is truthy (`&& value`, k8s.ts:295)            */
#define IGPU_ENTRY_SAME 0x08u   /* the same container holds the same key in its OTHER map with a
                                   strictly equal raw value: `req === lim` (PodsPage.tsx:69,
                                   PodDetailSection.tsx:80).  Strings never cross the boundary, so
                                   the marshal layer sets this bit on both entries of the pair.  */
#define IGPU_MAX_CONTAINERS 255 /* per list (containers / initContainers) of one pod           */
typedef struct igpu_entry {
  uint32_t pod;      /* row in the pod columns                                                 */
  uint16_t res;      /* resource dictionary id, < n_resources                                  */
  uint8_t flags;     /* IGPU_ENTRY_*                                                           */
  uint8_t container; /* index of the container in its list (spec.containers, or
                        spec.initContainers when IGPU_ENTRY_INIT), < IGPU_MAX_CONTAINERS      */
  int64_t value;     /* parseInt(v,10) || 0, |value| < 2^53                                    */
} igpu_entry;        /* 16 bytes, 16-byte aligned arrays                                       */

#define IGPU_OPT_POD_LIMITS 0x1u /* also produce the per-pod / per-node limits rollup           */
/* igpu_marshal_fleet_json only (never reaches igpu_fleet_cols.options): lay the pod columns out with the pods of
 * each node adjacent (stable, unscheduled pods last) and keep the permutation (igpu_marshal_pod_order).  Results of
 * a rollup do not depend on pod order; adjacent pods let the device add a node's pods up in a warp.               */
#define IGPU_OPT_GROUP_PODS_BY_NODE 0x100u

/* Input columns of one fleet (or one rank's node-range shard of it). All pointers are host
 * memory (pinned memory from igpu_host_alloc makes the copies asynchronous and fast).         */
typedef struct igpu_fleet_cols {
  uint32_t n_nodes;
  uint32_t n_pods;
  uint64_t n_entries;
  uint32_t n_resources; /* 4..IGPU_MAX_RESOURCES */
  uint32_t options;     /* IGPU_OPT_* */
  const uint8_t* node_flags;      /* [n_nodes] IGPU_NODE_*                                      */
  const int32_t* node_cap_i915;   /* [n_nodes] parseInt(capacity[i915]) || 0, 0 if absent       */
  const int32_t* node_cap_xe;     /* [n_nodes]                                                  */
  const int32_t* node_alloc_i915; /* [n_nodes] parseInt(allocatable[i915] ?? '0') counted only
                                     when capacity has the key (OverviewPage.tsx:96-99)          */
  const int32_t* node_alloc_xe;   /* [n_nodes]                                                  */
  const int32_t* node_alloc_own;  /* [n_nodes] sum over allocatable's own i915/xe keys of
                                     parseInt || 0 (NodeDetailSection.tsx:72-76)                 */
  const uint32_t* pod_node;       /* [n_pods] node row or IGPU_NODE_NONE                        */
  const uint8_t* pod_phase;       /* [n_pods] IGPU_PHASE_*                                      */
  const uint32_t* pod_entry_off;  /* [n_pods + 1] entries of pod p are [off[p], off[p+1])       */
  const igpu_entry* entries;      /* [n_entries]                                                */
} igpu_fleet_cols;

/* fleet_totals slots (int64).  Under multi-GPU these are the values after the allreduce.      */
enum {
  IGPU_T_GPU_NODES = 0,     /* filterIntelGpuNodes(...).length       k8s.ts:154-156            */
  IGPU_T_DISCRETE = 1,      /* OverviewPage.tsx:78-82                                          */
  IGPU_T_INTEGRATED = 2,
  IGPU_T_UNKNOWN = 3,
  IGPU_T_GPU_DEVICES = 4,   /* sum getNodeGpuCount                   OverviewPage.tsx:84        */
  IGPU_T_READY_NODES = 5,   /* OverviewPage.tsx:85                                             */
  IGPU_T_CAPACITY = 6,      /* OverviewPage.tsx:98  (NaN if slot IGPU_T_CAP_NAN != 0)          */
  IGPU_T_ALLOCATABLE = 7,   /* OverviewPage.tsx:99  (NaN if slot IGPU_T_ALLOC_NAN != 0)        */
  IGPU_T_CAP_NAN = 8,       /* number of GPU nodes whose capacity sum is NaN                   */
  IGPU_T_ALLOC_NAN = 9,
  IGPU_T_ALLOCATED = 10,    /* OverviewPage.tsx:105-113 (Running pods, i915 + xe requests)     */
  IGPU_T_GPU_PODS = 11,     /* filterGpuRequestingPods(...).length   k8s.ts:266-268            */
  IGPU_T_PHASE0 = 12,       /* 12..16: Running, Pending, Succeeded, Failed, Other (:122-130)   */
  IGPU_T_SCHEDULED_PODS = 17, /* GPU pods with a node (NodesPage.tsx:153-159)                  */
  IGPU_T_UTIL_PCT = 18,     /* Math.round(allocated / capacity * 100) or 0 (:115-116); filled
                               on the host side of the library after the allreduce             */
  IGPU_T_BAD_INPUT = 19,   /* pods / entries that broke the column invariants (offsets not ascending or past
                             n_entries, resource id >= n_resources, pod_node >= n_nodes, entries[i].pod not the
                             pod whose offsets cover i).  They are skipped on the device — never dereferenced —
                             and the call returns IGPU_ERR_ARG.                                            */
  IGPU_T_REQ0 = 32,         /* 32..63: fleet sum of Running pods' requests per resource id     */
  IGPU_T_LIM0 = 64,         /* 64..95: same for limits (IGPU_OPT_POD_LIMITS)                   */
  IGPU_TOTALS_LEN = 96
};

/* node_out_flags bits */
#define IGPU_NODEOUT_IS_GPU 0x01u    /* isIntelGpuNode, k8s.ts:125-152                          */
#define IGPU_NODEOUT_READY 0x02u     /* isNodeReady, k8s.ts:329-331                             */
#define IGPU_NODEOUT_CAP_NAN 0x04u   /* totalCapacity is NaN (NodesPage.tsx:173)                */
#define IGPU_NODEOUT_ALLOC_NAN 0x08u /* totalAllocatable is NaN (NodesPage.tsx:174)             */
#define IGPU_NODEOUT_BAR_DASH 0x10u  /* allocation bar shows '—' (NodesPage.tsx:36)             */

/* GpuType codes, k8s.ts:183-192 */
#define IGPU_TYPE_DISCRETE 0
#define IGPU_TYPE_INTEGRATED 1
#define IGPU_TYPE_UNKNOWN 2

/* Output columns. Any pointer may be NULL: that column is then not copied back. Planes are
 * column-major: plane r of pod_req is pod_req[r * n_pods + p].                                 */
typedef struct igpu_fleet_out {
  uint8_t* pod_is_gpu;      /* [n_pods] isGpuRequestingPod                k8s.ts:250-264        */
  uint32_t* pod_req_mask;   /* [n_pods] bit r set: key r is present in getPodGpuRequests(pod)   */
  int64_t* pod_req;         /* [n_resources][n_pods] getPodGpuRequests    k8s.ts:289-301        */
  uint32_t* pod_lim_mask;   /* [n_pods]              (IGPU_OPT_POD_LIMITS; no reference twin)   */
  int64_t* pod_lim;         /* [n_resources][n_pods] (IGPU_OPT_POD_LIMITS)                      */
  uint8_t* node_out_flags;  /* [n_nodes] IGPU_NODEOUT_*                                         */
  uint8_t* node_type;       /* [n_nodes] getNodeGpuType                   k8s.ts:185-192        */
  int64_t* node_gpu_count;  /* [n_nodes] getNodeGpuCount                  k8s.ts:171-180        */
  int64_t* node_cap_total;  /* [n_nodes] NodesPage.tsx:171-176 (see CAP_NAN flag)               */
  int64_t* node_alloc_total;/* [n_nodes]                                                        */
  uint32_t* node_pod_count; /* [n_nodes] podsByNode.get(name).length      NodesPage.tsx:153-159 */
  int64_t* node_bar_pct;    /* [n_nodes] GpuAllocationBar pct             NodesPage.tsx:35-38   */
  uint8_t* node_bar_level;  /* [n_nodes] 0 blue, 1 orange (>=70), 2 red (>=90)                  */
  int64_t* node_inuse;      /* [n_nodes] gpuInUse                  NodeDetailSection.tsx:77-87  */
  int64_t* node_util_pct;   /* [n_nodes] utilizationPct            NodeDetailSection.tsx:89     */
  uint8_t* node_util_level; /* [n_nodes] 0 success, 1 warning, 2 error  (:90-91)                */
  int64_t* node_req;        /* [n_resources][n_nodes] sum of pod_req over Running pods on node  */
  int64_t* node_lim;        /* [n_resources][n_nodes] (IGPU_OPT_POD_LIMITS)                     */
  int64_t* fleet_totals;    /* [IGPU_TOTALS_LEN]                                                */
  uint8_t* pod_gpu_containers; /* [n_pods] gpuContainers.length: MAIN containers with any prefixed key in
                               requests or limits, whatever its value (PodsPage.tsx:50-54,
                               PodDetailSection.tsx:48-52,106-108 "GPU Containers")                */
} igpu_fleet_out;

/* Timing and traffic of the most recent call on a context, from CUDA events on its stream.   */
typedef struct igpu_timing {
  float h2d_ms;
  float kernel_ms;
  float d2h_ms;
  float total_ms;
  uint64_t h2d_bytes;
  uint64_t d2h_bytes;
  uint64_t kernel_launches; /* cumulative count of this library's kernel launches on the ctx   */
  float kernel_detail_ms[12]; /* last launch of each kernel, IGPU_K_* below (0 = not run yet)    */
} igpu_timing;

/* indices into igpu_timing.kernel_detail_ms */
enum {
  IGPU_K_POD_ROLLUP = 0,
  IGPU_K_NODE_ROLLUP = 1,
  IGPU_K_WINDOW_STATS = 2,
  IGPU_K_QUANTILE_SAMPLE = 3,
  IGPU_K_QUANTILE_SCAN = 4,
  IGPU_K_QUANTILE_FINISH = 5,
  IGPU_K_QUANTILE_TAIL = 6, /* fallback + fleet reduction */
  IGPU_K_POWER_JOIN = 7,
  IGPU_K_SERIES_FUSED = 8,  /* igpu_series_rollup's single-read kernel */
  IGPU_K_COUNTER_RATE = 9   /* igpu_counters_rate */
};

/* ---- library / context --------------------------------------------------------------------- */

int32_t igpu_version(void); /* (major << 16) | (minor << 8) | patch */
const char* igpu_strerror(int32_t status);

/* device: CUDA ordinal.  stream: a cudaStream_t to launch on (e.g. the caller's current stream)
 * or NULL for a private non-blocking stream.                                                   */
int32_t igpu_ctx_create(int32_t device, void* stream, igpu_ctx** out);
int32_t igpu_ctx_destroy(igpu_ctx* ctx);
const char* igpu_last_error(const igpu_ctx* ctx);
int32_t igpu_ctx_sync(igpu_ctx* ctx);
/* The IGPU_* environment switches (kernel variants for comparison, sweep knobs; listed in csrc/common.cuh) are read when a
 * context is created, never on a call path; this reads them again.  No counterpart in the reference.                  */
int32_t igpu_ctx_reload_env(igpu_ctx* ctx);
int32_t igpu_get_timing(const igpu_ctx* ctx, igpu_timing* out);

/* Pinned host memory for callers that want asynchronous, full-rate PCIe copies.              */
int32_t igpu_host_alloc(uint64_t bytes, void** out);
int32_t igpu_host_free(void* p);

/* Optional communicator: one rank per GPU.  Rank 0 obtains a 128-byte id, the host program
 * distributes it (any transport), every rank then calls igpu_comm_init.  With a communicator
 * attached the fleet totals and the fleet power totals are all-reduced over NCCL on the
 * context's stream right after the kernel that produces them.  NCCL is located with dlopen at
 * this point; a process that never calls these needs no NCCL.  The reference has no
 * counterpart (single browser thread).                                                        */
int32_t igpu_comm_unique_id(void* out128);
int32_t igpu_comm_init(igpu_ctx* ctx, const void* id128, int32_t rank, int32_t nranks);
int32_t igpu_comm_destroy(igpu_ctx* ctx);

/* ---- allocation rollup + classification (K1) -------------------------------------------------
 * Replaces, for a whole fleet at once:
 *   isIntelGpuNode/filterIntelGpuNodes k8s.ts:125-156      getNodeGpuType   k8s.ts:185-192
 *   getNodeGpuCount   k8s.ts:171-180                        isNodeReady      k8s.ts:329-331
 *   isGpuRequestingPod/filterGpuRequestingPods k8s.ts:250-268
 *   getPodGpuRequests k8s.ts:289-301
 *   Overview loops    src/components/OverviewPage.tsx:71-130
 *   Nodes page loops  src/components/NodesPage.tsx:35-38,152-189,270-273
 *   Node detail loops src/components/NodeDetailSection.tsx:61-91                                */

/* One-shot call with host buffers: upload, reduce, download, synchronise.                      */
int32_t igpu_alloc_rollup(igpu_ctx* ctx, const igpu_fleet_cols* cols, const igpu_fleet_out* out);

/* Staged form: device-resident columns that can be reduced repeatedly.                         */
int32_t igpu_fleet_create(igpu_ctx* ctx, uint32_t max_nodes, uint32_t max_pods, uint64_t max_entries,
                          uint32_t n_resources, uint32_t options, igpu_fleet** out);
int32_t igpu_fleet_upload(igpu_ctx* ctx, igpu_fleet* fleet, const igpu_fleet_cols* cols); /* async */
int32_t igpu_fleet_rollup(igpu_ctx* ctx, igpu_fleet* fleet);                               /* async */
int32_t igpu_fleet_download(igpu_ctx* ctx, igpu_fleet* fleet, const igpu_fleet_out* out);  /* syncs */
int32_t igpu_fleet_destroy(igpu_ctx* ctx, igpu_fleet* fleet);

/* Per-container key sets of the GPU containers (the filter and the Set of PodsPage.tsx:49-62 and
 * PodDetailSection.tsx:48-65, for every pod of the fleet at once).  One row per MAIN container
 * that has at least one prefixed key, in (pod, container) order; bit r of a mask speaks about
 * resource id r.  What the two views print follows from the masks:
 *   PodsPage.tsx:69-73      collapsed "key: v"          = req_truthy & lim_truthy & same
 *                           "key: req=.. lim=.."         = (req_truthy | lim_truthy) & ~collapsed
 *                           nothing                      = neither value truthy
 *   PodDetailSection.tsx:71-83  request row for every key of req_keys | lim_keys;
 *                           limit row                    = lim_truthy & ~same
 * The order of the keys inside a container is the order of its entries (first occurrence).     */
typedef struct igpu_container_row {
  uint32_t pod;        /* row in the pod columns                                                */
  uint32_t container;  /* index in spec.containers                                              */
  uint32_t req_keys;   /* keys present in requests                                              */
  uint32_t lim_keys;   /* keys present in limits                                                */
  uint32_t req_truthy; /* ... whose request value is truthy                                     */
  uint32_t lim_truthy; /* ... whose limit value is truthy                                       */
  uint32_t same;       /* keys in both maps with req === lim                                    */
  uint32_t n_entries;  /* entries of this container (requests + limits)                         */
} igpu_container_row;  /* 32 bytes                                                              */
/* After igpu_fleet_rollup.  Writes min(n_rows, capacity) rows to host memory and the true row count to
 * *n_rows (call with capacity 0 to size the buffer; never more than n_entries rows).  Synchronises. */
int32_t igpu_fleet_container_rows(igpu_ctx* ctx, igpu_fleet* fleet, igpu_container_row* host_rows, uint64_t capacity,
                                  uint64_t* n_rows);

/* ---- native host marshal: Kubernetes JSON text -> igpu_fleet_cols --------------------------------
 * Replaces JSON.parse + the per-object field access that feeds every helper of k8s.ts (labels and
 * capacity k8s.ts:128-149, conditions :330, containers / initContainers resources :254-261,:291-297,
 * spec.nodeName NodesPage.tsx:155, status.phase OverviewPage.tsx:106,123) and extractJsonData
 * (IntelGpuDataContext.tsx:85-90).  Each text is a JSON array of objects or a List object
 * {"items":[...]} (k8s.ts:315-323).  The JS semantics (parseInt(v,10) || 0, truthiness, === 'true',
 * ?? '0') are applied here, on `threads` host threads (<= 0: all cores).  Host code only: it works
 * without a GPU.  Errors: malformed JSON -> IGPU_ERR_ARG, quantities beyond 2^53 / 32 resource names
 * -> IGPU_ERR_RANGE; text via igpu_last_error(NULL).                                              */
typedef struct igpu_marshal igpu_marshal;
int32_t igpu_marshal_fleet_json(const char* nodes_json, uint64_t nodes_len, const char* pods_json, uint64_t pods_len,
                                uint32_t options, int32_t threads, igpu_marshal** out);
/* Fills `cols` with pointers into the marshal object (valid until igpu_marshal_free). */
int32_t igpu_marshal_columns(const igpu_marshal* m, igpu_fleet_cols* cols);
/* Resource dictionary: name of id, NULL beyond the last id (ids 0..4 are the five known names). */
/* [n_pods]: order[i] = position, in the pods JSON, of the pod in column position i; NULL unless
 * IGPU_OPT_GROUP_PODS_BY_NODE was given.                                                        */
const uint32_t* igpu_marshal_pod_order(const igpu_marshal* m);
const char* igpu_marshal_resource_name(const igpu_marshal* m, uint32_t id);
int32_t igpu_marshal_free(igpu_marshal* m);

/* ---- instant power join + fleet sums (K2) ----------------------------------------------------
 * Replaces fetchGpuMetrics' join, src/api/metrics.ts:118-149, and the fleet reduce at
 * src/components/MetricsPage.tsx:330-332 with PowerBar :50-53.  chip / instance / node names are
 * dictionary ids.  Value columns hold parseFloat(value[1]) (metrics.ts:130,137).
 * The maps are keyed by chip id only and the LAST row for a chip id wins (Map.set).           */
typedef struct igpu_power_cols {
  uint32_t n_chips, n_energy, n_power_max, n_uname;
  uint32_t n_chip_ids;      /* dictionary sizes */
  uint32_t n_instance_ids;
  const uint32_t* chip_chip;        /* [n_chips] chip id of each chip-identification row        */
  const uint32_t* chip_instance;    /* [n_chips] instance id                                    */
  const uint32_t* energy_chip;      /* [n_energy] chip id, IGPU_NODE_NONE when the label is falsy */
  const double* energy_value;       /* [n_energy]                                               */
  const uint32_t* power_max_chip;   /* [n_power_max]                                            */
  const double* power_max_value;    /* [n_power_max]                                            */
  const uint32_t* uname_instance;   /* [n_uname] instance id, IGPU_NODE_NONE when falsy         */
  const uint32_t* uname_node;       /* [n_uname] node-name id after the ?? chain (metrics.ts:122) */
} igpu_power_cols;

#define IGPU_POWER_HAS_WATTS 0x1u
#define IGPU_POWER_HAS_MAX 0x2u
#define IGPU_POWER_HAS_NODE 0x4u
typedef struct igpu_power_out {
  uint8_t* chip_flags;      /* [n_chips] IGPU_POWER_HAS_*: null-ness of powerWatts / powerMaxWatts,
                               and whether instanceToNode had the instance (else nodeName=instance) */
  double* chip_watts;       /* [n_chips] powerWatts (undefined where the flag is clear)          */
  double* chip_max_watts;   /* [n_chips]                                                        */
  uint32_t* chip_node;      /* [n_chips] node-name id                                           */
  double* fleet;            /* [4] sum watts, sum max watts, pct (NaN = null), n chips          */
} igpu_power_out;
int32_t igpu_power_join_sum(igpu_ctx* ctx, const igpu_power_cols* cols, const igpu_power_out* out);

/* The same join and fleet sums for callers that must not allocate.  It has no counterpart in the
 * reference: it exists for host programs that cannot let this library call cudaMalloc — a server
 * that reserved all device memory for its own pool at start-up (a later cudaMalloc just fails), a
 * scrape loop that shares the GPU with other streams and cannot afford the device-wide
 * synchronisation hidden in cudaMalloc / cudaFree, or a program whose policy is that only its own
 * allocator touches device memory.
 *  - Every pointer inside `cols` and `out` is a DEVICE pointer here (the two structs themselves are
 *    host memory); all five outputs are required.  out->fleet receives 4 doubles:
 *    sum watts, sum max watts, PowerBar percentage (NaN = none), number of chips.
 *  - No context is needed and nothing is allocated, freed or copied to the host: the intermediate
 *    tables live in a fixed array inside the library image (3 MB per GPU that loads the library, held
 *    until the process exits).  That fixes the limit: at most IGPU_STATIC_JOIN_MAX_IDS distinct chip
 *    ids and as many instance ids; larger dictionaries are refused with IGPU_ERR_RANGE (text via
 *    igpu_last_error(NULL)) — use igpu_power_join_sum for those.
 *  - The work is queued on `stream` (a cudaStream_t, NULL = the default stream) and the call returns
 *    without waiting; synchronise the stream before reading the outputs.
 *  - The first call on a device does a one-time set-up (it looks up the device address of the
 *    built-in array); make that first call before you start any stream capture.  With CUDA's lazy
 *    module loading that look-up is also the moment the 3 MB array gets its device memory, so a
 *    program that reserves all of device memory for its own pool must make this first call BEFORE
 *    it does so; afterwards the set-up can fail (IGPU_ERR_CUDA) and nothing is launched.
 *  - `device` must be an ordinal this machine has (checked against cudaGetDeviceCount); an unknown
 *    ordinal is IGPU_ERR_ARG and touches no device.
 *  - All calls on a device share the one array, which keeps what the previous call left in it; each
 *    call clears the part it reads.  Two calls whose GPU work overlaps (different streams, or
 *    different threads) race on it — run them one after another.                                  */
#define IGPU_STATIC_JOIN_MAX_IDS (1u << 18)
int32_t igpu_power_join_sum_static(int32_t device, void* stream, const igpu_power_cols* dev_cols,
                                   const igpu_power_out* dev_out);

/* ---- power time series (K3 window statistics, K4 per-node quantile) ---------------------------
 * No reference counterpart exists: metrics.ts issues instant queries only (:67-75,:101-116) and
 * the only windowed arithmetic, rate(...[5m]) (:106), runs inside Prometheus.  Semantics follow
 * PromQL's min/avg/max_over_time and quantile_over_time as fixed in SURVEY.md §8(c):
 *  - layout [series][time] float32 on a regular grid, series grouped by node;
 *  - NaN = missing sample, ignored by every statistic and by the counts;
 *  - window w of a series covers samples [w*W, min((w+1)*W, T)); min/max are exact float32 order
 *    statistics, mean = (float64 sum)/count rounded to float32; an empty window yields NaN;
 *  - quantile mode 0 (linear): rank = q*(n-1), v = x[lo]*(1-w) + x[hi]*w in float64 over the
 *    node's sorted valid samples; mode 1 (nearest rank): x[ceil(q*n)-1]; n == 0 yields NaN.    */
typedef struct igpu_node_series_stats {
  double sum;
  uint64_t count; /* valid (non-NaN) samples */
  float min;
  float max;
} igpu_node_series_stats; /* 24 bytes */

/* fleet vector written by the quantile call (after the allreduce when a communicator is attached) */
enum { IGPU_F_SUM = 0, IGPU_F_COUNT = 1, IGPU_F_MIN = 2, IGPU_F_MAX = 3, IGPU_F_NODES = 4, IGPU_FLEET_LEN = 8 };

/* Rows may have any length: on the device they are stored at a pitch of a multiple of 4 samples (the padding is NaN,
 * i.e. no sample, and never counted), so every kernel sees 16-byte-aligned rows; upload / download / generate and all
 * results speak about n_time.                                                                                      */
int32_t igpu_series_create(igpu_ctx* ctx, uint32_t n_series, uint32_t n_time, uint32_t n_nodes,
                           const uint32_t* node_series_off /* [n_nodes+1], ascending, host */,
                           igpu_series** out);
int32_t igpu_series_destroy(igpu_ctx* ctx, igpu_series* s);
/* Copy rows [first, first+count) from / to host memory (async upload, synchronous download).   */
int32_t igpu_series_upload(igpu_ctx* ctx, igpu_series* s, uint32_t first, uint32_t count, const float* host);
int32_t igpu_series_download(igpu_ctx* ctx, igpu_series* s, uint32_t first, uint32_t count, float* host);
/* Fill the slab on the device with the synthetic fleet of DESIGN.md §generator; global_first is
 * the fleet-wide index of row 0 (so shards of one fleet generate disjoint, reproducible rows);
 * nan_per_million > 0 knocks out that share of samples as NaN.                                 */
int32_t igpu_series_generate(igpu_ctx* ctx, igpu_series* s, uint64_t seed, uint64_t global_first,
                             uint32_t nan_per_million);

/* Window statistics over every series. Results stay on the device; non-NULL host pointers
 * ([n_series][ceil(T/W)] each) also receive a copy. */
int32_t igpu_series_window_stats(igpu_ctx* ctx, igpu_series* s, uint32_t window, float* host_min,
                                 float* host_mean, float* host_max);
int32_t igpu_series_read_window_stats(igpu_ctx* ctx, igpu_series* s, uint32_t first, uint32_t count,
                                      float* host_min, float* host_mean, float* host_max);

/* Per-node quantile (p95 with q = 0.95) + per-node sum/count/min/max + fleet totals.           */
int32_t igpu_series_quantile_by_node(igpu_ctx* ctx, igpu_series* s, double q, int32_t mode,
                                     double* host_quantile /* [n_nodes] */,
                                     igpu_node_series_stats* host_stats /* [n_nodes] or NULL */,
                                     double* host_fleet /* [IGPU_FLEET_LEN] or NULL */);

/* Window statistics AND per-node quantile in ONE read of the slab (same results as the two calls
 * above, same host-pointer conventions; NULL host pointers leave that result on the device).  This is
 * the call a range view makes (docs/screenshots/03-metrics.svg:26-37 mocks "Peak / Avg power" over a
 * time range; the reference never implemented it).  Shapes the single-read kernel does not cover fall
 * back to the two separate passes.                                                              */
int32_t igpu_series_rollup(igpu_ctx* ctx, igpu_series* s, uint32_t window, double q, int32_t mode,
                           float* host_min, float* host_mean, float* host_max,
                           double* host_quantile /* [n_nodes] or NULL */,
                           igpu_node_series_stats* host_stats /* [n_nodes] or NULL */,
                           double* host_fleet /* [IGPU_FLEET_LEN] or NULL */);

/* Fleet-level views of the slab (north star: "per-node and per-fleet windowed min/mean/max/p95").  The reference
 * has no counterpart (only the mock-up docs/screenshots/03-metrics.svg:26-37,85-89 shows a range view); the
 * definitions are the PromQL ones of SURVEY.md §8(c) (the CPU oracle restates them).  With a communicator
 * attached the results speak about the WHOLE fleet: every rank passes its node-range shard and gets the same answer.
 *
 * igpu_series_quantile_fleet: the quantile over every valid sample of every series (same rank / interpolation
 * rules as igpu_series_quantile_by_node; NaN when there is no valid sample).  Exact, in ONE read of the slab where the
 * fleet is large enough to sample: (1) a radix select over a sparse sample (one 128-byte line in 64) of every rank's
 * slab brackets the wanted ranks by two keys, 5 standard deviations of the sample rank either side, a line counted as one
 * draw; (2) one pass counts the valid samples and those below the bracket and gathers the samples inside it (about a
 * thousandth of the slab); (3) a radix select over the gathered samples gives the exact ranks.  Every histogram and the
 * counters are all-reduced (int64 sum), so all ranks take the same decisions.  The counts of (2) prove whether the
 * bracket held the ranks; if not (or the fleet is small) the select runs over the whole slab in three reads.       */
int32_t igpu_series_quantile_fleet(igpu_ctx* ctx, igpu_series* s, double q, int32_t mode, double* host_quantile /* [1] */,
                                   uint64_t* host_count /* [1] valid samples in the fleet, or NULL */);

/* Window statistics over a node's chips and over the fleet: min / max of the samples, mean = sum of the samples
 * / number of valid samples, count = that number; NaN (count 0) where a window holds no valid sample.  Node planes
 * are [n_nodes][ceil(T/W)] (this rank's nodes), fleet planes [ceil(T/W)] (all-reduced).  Any pointer may be NULL.
 * One read of the slab — or none: after igpu_series_keep_window_counts(s, 1), an igpu_series_rollup of the same
 * window that ran its single-read kernel leaves the per-series planes and valid-sample counts on the device, and
 * this call folds those over each node's chips instead (min of mins, max of maxes, sum of counts, count-weighted
 * mean of the float32 means: min / max / count exact, mean within float32 rounding of the per-series means, i.e. a
 * relative 6e-8).  Any upload / generate / rate into the slab in between, or another window, brings the read back. */
/* Makes igpu_series_rollup keep the number of valid samples per (series, window) next to its min / mean / max planes
 * (uint16: windows of up to 65,535 samples; 2 bytes per series-window more written by the single-read kernel).  Off by
 * default.  No counterpart in the reference.                                                                        */
/* full passes over the slab the last igpu_series_quantile_fleet on s made: 1 bracketed, 3 exact, 4 when the bracket missed */
uint32_t igpu_series_fleet_quantile_reads(const igpu_series* s);
int32_t igpu_series_keep_window_counts(igpu_ctx* ctx, igpu_series* s, int32_t on);
typedef struct igpu_window_rollup_out {
  float* node_min;
  float* node_mean;
  float* node_max;
  uint32_t* node_count;
  float* fleet_min;
  float* fleet_mean;
  float* fleet_max;
  uint64_t* fleet_count;
  uint32_t* slab_reads; /* [1] passes over the slab this call made: 1, or 0 (see igpu_series_keep_window_counts) */
} igpu_window_rollup_out;
int32_t igpu_series_window_rollup(igpu_ctx* ctx, igpu_series* s, uint32_t window, const igpu_window_rollup_out* out);

/* ---- N3: rate() over cumulative energy counters, on the device --------------------------------
 * Replaces the server-side `rate(node_hwmon_energy_joule_total[5m])` of metrics.ts:106-109 when the
 * plugin fetches the raw counter as a `query_range` matrix instead of four instant queries
 * (SURVEY.md §8(f) N3).  The algorithm is Prometheus' extrapolatedRate (promql/functions.go, v3.x:
 * left-open range, counter-reset correction, zero-point clamp, 1.1x extrapolation threshold); it is a
 * third-party server, not part of the reference repository, so parity is pinned only by the promql
 * known answers restated in tests/ ("parity unpinned by the reference").
 *
 * Layout: float64 [n_series][n_time] on a regular time grid t_i = i * step_ms (what query_range
 * returns); NaN = no sample at that step.  Evaluation point e looks at grid samples
 * (e - range_samples, e] (or [e - range_samples, e] with IGPU_RATE_CLOSED_LEFT, Prometheus 2.x).
 * Output j = rate at e = first_eval + j * eval_stride, as float32 into an igpu_series slab with the
 * same series count and n_time = igpu_counters_rate_len(...); fewer than two samples in range -> NaN.  */
typedef struct igpu_counters igpu_counters;

enum {
  IGPU_RATE_CLOSED_LEFT = 1u, /* Prometheus 2.x range selector [t-R, t]; default is 3.x (t-R, t]     */
  IGPU_RATE_INCREASE = 2u,    /* increase(): do not divide by the range                             */
  IGPU_RATE_GAUGE = 4u        /* delta(): no counter-reset correction, no zero-point clamp          */
};

typedef struct igpu_rate_params {
  uint32_t step_ms;       /* grid step (scrape / query_range step), > 0                            */
  uint32_t range_samples; /* range = range_samples * step_ms; 5m at a 15 s step = 20; >= 1         */
  uint32_t first_eval;    /* grid index of output 0                                                */
  uint32_t eval_stride;   /* grid steps between outputs, >= 1                                      */
  uint32_t options;       /* IGPU_RATE_*                                                           */
} igpu_rate_params;

int32_t igpu_counters_create(igpu_ctx* ctx, uint32_t n_series, uint32_t n_time, igpu_counters** out);
int32_t igpu_counters_destroy(igpu_ctx* ctx, igpu_counters* c);
int32_t igpu_counters_upload(igpu_ctx* ctx, igpu_counters* c, uint32_t first, uint32_t count, const double* host);
/* Synthetic energy counters (joules): per-series constant power 50-300 W plus bounded jitter,
 * a counter reset every reset_period samples (0 = never) at a per-series phase, missing_per_million
 * samples knocked out as NaN.  Integer arithmetic + one rounding, so the oracle reproduces it bit
 * for bit.                                                                                       */
int32_t igpu_counters_generate(igpu_ctx* ctx, igpu_counters* c, uint64_t seed, uint64_t global_first,
                               uint32_t step_ms, uint32_t reset_period, uint32_t missing_per_million);
/* The slab's gap statistic: upload and generate look at a sparse sample of what they wrote (one 128-byte line in 64;
 * small slabs whole) and keep (missing, looked at).  igpu_counters_rate has two kernels for ranges of up to 33
 * samples — one that streams and handles gaps block by block out of line, one that takes every block from the
 * presence masks — and the device picks by this statistic (more than 3,000 missing per million -> the second); both
 * give the same bits.  Synchronises the stream.  No counterpart in the reference.                  */
int32_t igpu_counters_gap_stat(igpu_ctx* ctx, const igpu_counters* c, uint64_t* missing, uint64_t* looked_at);
/* number of outputs per series for a slab of n_time samples (0 when first_eval >= n_time)        */
uint32_t igpu_counters_rate_len(uint32_t n_time, const igpu_rate_params* p);
/* dst: same n_series, n_time == igpu_counters_rate_len(c.n_time, p); stays on the device, ready for
 * igpu_series_rollup.  host_rate64 (optional, [n_series][rate_len] doubles) receives the unrounded
 * float64 results (test hook; costs a second output plane).
 * Every range length is bit-identical to the serial evaluation.  Ranges of up to 33 samples stream through one of
 * two kernels (see igpu_counters_gap_stat); longer ranges through a third that keeps the last 256 samples of history in
 * shared memory (beyond that: reads the first sample of a range from global memory) and leaves blocks with a gap, a
 * reset or a clamp candidate in reach to a follow-up kernel.  n_time < 2^31.                                        */
int32_t igpu_counters_rate(igpu_ctx* ctx, const igpu_counters* c, const igpu_rate_params* p, igpu_series* dst,
                           double* host_rate64);

/* Prometheus `query_range` response (HTTP API v1 "matrix") -> float64 [series][time] slab on the host, ready for
 * igpu_counters_upload.  Replaces JSON.parse + the parseFloat loops of metrics.ts:67-75,130-141 for the range view.
 * Grid: t = start_ms + i * step_ms, i < n_time; steps without a sample are NaN; samples off the grid or outside it
 * are skipped and counted (igpu_matrix_dropped).  `status` other than "success" gives zero series (metrics.ts:73).
 * A sample's text is read the way the reference reads it, with the rules of JS parseFloat (metrics.ts:130-137): longest
 * decimal-literal prefix, so "NaN", "+Inf" and "-Inf" (Prometheus' spellings) all come out NaN = no sample.             
 * label_names: the metric labels to keep per series (e.g. "chip", "instance"); CPU only, multithreaded.          */
typedef struct igpu_matrix igpu_matrix;
int32_t igpu_marshal_matrix_json(const char* json, uint64_t len, int64_t start_ms, uint32_t step_ms, uint32_t n_time,
                                 const char* const* label_names, uint32_t n_labels, int32_t threads, igpu_matrix** out);
uint32_t igpu_matrix_n_series(const igpu_matrix* m);
const double* igpu_matrix_values(const igpu_matrix* m); /* [n_series][n_time] */
const char* igpu_matrix_label(const igpu_matrix* m, uint32_t series, uint32_t label); /* NULL = label absent */
uint64_t igpu_matrix_dropped(const igpu_matrix* m);
int32_t igpu_matrix_free(igpu_matrix* m);

#ifdef __cplusplus
}
#endif
#endif /* IGPU_ROLLUP_H */
