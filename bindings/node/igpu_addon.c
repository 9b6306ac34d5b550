/* igpu_addon.c — Node-API addon over libigpu_rollup.so.  There is no Node.js in this image:
 * the code below is EXECUTED by the test suite against a stand-in for the N-API calls it uses (tests/napi_shim/, a toy
 * value model in C driven from pytest), which checks its marshalling, its argument validation and its results against the
 * ctypes path; under a real Node.js it is unverified.  Build against a real Node.js with
 *   cc -shared -fPIC -DIGPU_HAVE_NODE_API_H -I<node>/include/node -I../../include igpu_addon.c \
 *      -L../../headlamp_intel_gpu_plugin_b200/csrc -ligpu_rollup -o igpu_addon.node
 *
 * Exports (all synchronous; the TypeScript layer wraps the batch calls in a worker when needed):
 *   rollupFleet(cols)  -> outputs     igpu_alloc_rollup   (k8s.ts:125-301 + the page loops)
 *   rollupFleetJson(nodesJson, podsJson) -> outputs + resourceNames
 *                                     igpu_marshal_fleet_json + igpu_alloc_rollup: the two List responses as
 *                                     UTF-8 text (Buffer / Uint8Array), no JSON.parse on the JS side
 *   powerJoinSum(cols) -> outputs     igpu_power_join_sum (metrics.ts:118-149, MetricsPage.tsx:330-332)
 *   rangeView(series)  -> outputs     igpu_series_rollup (+ igpu_series_quantile_fleet, igpu_series_window_rollup): the
 *                                     range view the plugin only mocks up (docs/screenshots/03-metrics.svg:26-45,85-89)
 * Typed arrays are passed zero-copy; outputs are fresh typed arrays owned by V8.
 * There is no CPU fallback: if no B200 context can be created the module throws on load and the
 * TypeScript layer keeps using the reference path (which is all a browser bundle can do anyway). */
#ifdef IGPU_HAVE_NODE_API_H
#include <node_api.h>
#else
#include "node_api_min.h"
#endif
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/igpu_rollup.h"

static igpu_ctx* g_ctx;

#define CHECK(call)                                                   \
  do {                                                                \
    if ((call) != napi_ok) {                                          \
      napi_throw_error(env, NULL, "igpu addon: N-API call failed: " #call); \
      return NULL;                                                    \
    }                                                                 \
  } while (0)

/* The data pointer of the typed array `obj[name]`, which must be of type `want` and, when want_len != SIZE_MAX, hold exactly
 * want_len elements.  On any mismatch *bad is set and a TypeError-style exception names the column: the library would
 * otherwise read past a V8 buffer. */
static void* ta_col(napi_env env, napi_value obj, const char* name, napi_typedarray_type want, size_t want_len, size_t* len, int* bad) {
  napi_value v, ab;
  void* data = NULL;
  napi_typedarray_type t;
  size_t off, n = 0;
  char msg[160];
  if (*bad) return NULL;
  if (napi_get_named_property(env, obj, name, &v) != napi_ok || napi_get_typedarray_info(env, v, &t, &n, &data, &ab, &off) != napi_ok) {
    snprintf(msg, sizeof msg, "igpu addon: column '%s' is missing or not a typed array", name);
  } else if (t != want) {
    snprintf(msg, sizeof msg, "igpu addon: column '%s' has the wrong element type", name);
  } else if (want_len != (size_t)-1 && n != want_len) {
    snprintf(msg, sizeof msg, "igpu addon: column '%s' has %zu elements, expected %zu", name, n, want_len);
  } else {
    if (len) *len = n;
    return data;
  }
  *bad = 1;
  napi_throw_error(env, "ERR_INVALID_ARG_TYPE", msg);
  return NULL;
}
#define ANY_LEN ((size_t)-1)

static napi_value new_ta(napi_env env, napi_value out, const char* name, napi_typedarray_type type, size_t count, size_t elem, void** data) {
  napi_value ab, ta;
  if (napi_create_arraybuffer(env, (count * elem) > 0 ? count * elem : 1, data, &ab) != napi_ok) return NULL;
  if (napi_create_typedarray(env, type, count, ab, 0, &ta) != napi_ok) return NULL;
  napi_set_named_property(env, out, name, ta);
  return ta;
}

static napi_value run_rollup(napi_env env, const igpu_fleet_cols* c);

/* rollupFleet({nodeFlags:Uint8Array, capI915:Int32Array, capXe, allocI915, allocXe, allocOwn,
 *              podNode:Uint32Array, podPhase:Uint8Array, podEntryOff:Uint32Array, entries:Uint8Array(16 B records),
 *              nResources:number, options:number}) */
static napi_value RollupFleet(napi_env env, napi_callback_info info) {
  size_t argc = 1, n_nodes = 0, n_pods = 0, n_off = 0, n_bytes = 0;
  int bad = 0;
  napi_value arg, out, v;
  uint32_t R = 5, options = IGPU_OPT_POD_LIMITS;
  igpu_fleet_cols c;
  igpu_fleet_out o;
  CHECK(napi_get_cb_info(env, info, &argc, &arg, NULL, NULL));
  memset(&c, 0, sizeof c);
  memset(&o, 0, sizeof o);
  c.node_flags = (const uint8_t*)ta_col(env, arg, "nodeFlags", napi_uint8_array, ANY_LEN, &n_nodes, &bad);
  c.node_cap_i915 = (const int32_t*)ta_col(env, arg, "capI915", napi_int32_array, n_nodes, NULL, &bad);
  c.node_cap_xe = (const int32_t*)ta_col(env, arg, "capXe", napi_int32_array, n_nodes, NULL, &bad);
  c.node_alloc_i915 = (const int32_t*)ta_col(env, arg, "allocI915", napi_int32_array, n_nodes, NULL, &bad);
  c.node_alloc_xe = (const int32_t*)ta_col(env, arg, "allocXe", napi_int32_array, n_nodes, NULL, &bad);
  c.node_alloc_own = (const int32_t*)ta_col(env, arg, "allocOwn", napi_int32_array, n_nodes, NULL, &bad);
  c.pod_node = (const uint32_t*)ta_col(env, arg, "podNode", napi_uint32_array, ANY_LEN, &n_pods, &bad);
  c.pod_phase = (const uint8_t*)ta_col(env, arg, "podPhase", napi_uint8_array, n_pods, NULL, &bad);
  c.pod_entry_off = (const uint32_t*)ta_col(env, arg, "podEntryOff", napi_uint32_array, n_pods + 1, &n_off, &bad);
  c.entries = (const igpu_entry*)ta_col(env, arg, "entries", napi_uint8_array, ANY_LEN, &n_bytes, &bad);
  if (bad) return NULL;
  if (n_bytes % sizeof(igpu_entry) || (n_off && c.pod_entry_off[n_off - 1] != n_bytes / sizeof(igpu_entry))) {
    napi_throw_error(env, "ERR_INVALID_ARG_VALUE", "igpu addon: 'entries' must hold podEntryOff[nPods] records of 16 bytes");
    return NULL;
  }
  if (napi_get_named_property(env, arg, "nResources", &v) == napi_ok) napi_get_value_uint32(env, v, &R);
  if (napi_get_named_property(env, arg, "options", &v) == napi_ok) napi_get_value_uint32(env, v, &options);
  c.n_nodes = (uint32_t)n_nodes;
  c.n_pods = (uint32_t)n_pods;
  c.n_entries = n_bytes / sizeof(igpu_entry);
  c.n_resources = R;
  c.options = options;
  (void)out;
  (void)o;
  return run_rollup(env, &c);
}

/* shared by rollupFleet and rollupFleetJson: allocate the output typed arrays and run the rollup */
static napi_value run_rollup(napi_env env, const igpu_fleet_cols* c) {
  napi_value out;
  igpu_fleet_out o;
  const size_t n_nodes = c->n_nodes, n_pods = c->n_pods, R = c->n_resources;
  memset(&o, 0, sizeof o);
  CHECK(napi_create_object(env, &out));
  new_ta(env, out, "podIsGpu", napi_uint8_array, n_pods, 1, (void**)&o.pod_is_gpu);
  new_ta(env, out, "podGpuContainers", napi_uint8_array, n_pods, 1, (void**)&o.pod_gpu_containers);
  new_ta(env, out, "podReqMask", napi_uint32_array, n_pods, 4, (void**)&o.pod_req_mask);
  new_ta(env, out, "podReq", napi_bigint64_array, R * n_pods, 8, (void**)&o.pod_req);
  if (c->options & IGPU_OPT_POD_LIMITS) {
    new_ta(env, out, "podLimMask", napi_uint32_array, n_pods, 4, (void**)&o.pod_lim_mask);
    new_ta(env, out, "podLim", napi_bigint64_array, R * n_pods, 8, (void**)&o.pod_lim);
    new_ta(env, out, "nodeLim", napi_bigint64_array, R * n_nodes, 8, (void**)&o.node_lim);
  }
  new_ta(env, out, "nodeOutFlags", napi_uint8_array, n_nodes, 1, (void**)&o.node_out_flags);
  new_ta(env, out, "nodeType", napi_uint8_array, n_nodes, 1, (void**)&o.node_type);
  new_ta(env, out, "nodeGpuCount", napi_bigint64_array, n_nodes, 8, (void**)&o.node_gpu_count);
  new_ta(env, out, "nodeCapTotal", napi_bigint64_array, n_nodes, 8, (void**)&o.node_cap_total);
  new_ta(env, out, "nodeAllocTotal", napi_bigint64_array, n_nodes, 8, (void**)&o.node_alloc_total);
  new_ta(env, out, "nodePodCount", napi_uint32_array, n_nodes, 4, (void**)&o.node_pod_count);
  new_ta(env, out, "nodeBarPct", napi_bigint64_array, n_nodes, 8, (void**)&o.node_bar_pct);
  new_ta(env, out, "nodeBarLevel", napi_uint8_array, n_nodes, 1, (void**)&o.node_bar_level);
  new_ta(env, out, "nodeInUse", napi_bigint64_array, n_nodes, 8, (void**)&o.node_inuse);
  new_ta(env, out, "nodeUtilPct", napi_bigint64_array, n_nodes, 8, (void**)&o.node_util_pct);
  new_ta(env, out, "nodeUtilLevel", napi_uint8_array, n_nodes, 1, (void**)&o.node_util_level);
  new_ta(env, out, "nodeReq", napi_bigint64_array, R * n_nodes, 8, (void**)&o.node_req);
  new_ta(env, out, "fleetTotals", napi_bigint64_array, IGPU_TOTALS_LEN, 8, (void**)&o.fleet_totals);
  if (igpu_alloc_rollup(g_ctx, c, &o) != IGPU_OK) {
    napi_throw_error(env, NULL, igpu_last_error(g_ctx));
    return NULL;
  }
  return out;
}

/* rollupFleetJson(nodesJson: Uint8Array, podsJson: Uint8Array): the List responses as UTF-8 text */
static napi_value RollupFleetJson(napi_env env, napi_callback_info info) {
  size_t argc = 2, nlen = 0, plen = 0, off;
  napi_value argv[2], ab, out, names, v;
  napi_typedarray_type t;
  void *ndata = NULL, *pdata = NULL;
  igpu_marshal* m = NULL;
  igpu_fleet_cols c;
  uint32_t i;
  CHECK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  CHECK(napi_get_typedarray_info(env, argv[0], &t, &nlen, &ndata, &ab, &off));
  CHECK(napi_get_typedarray_info(env, argv[1], &t, &plen, &pdata, &ab, &off));
  /* pods come out in the order of the JSON (no IGPU_OPT_GROUP_PODS_BY_NODE): the TypeScript side indexes by position */
  if (igpu_marshal_fleet_json((const char*)ndata, nlen, (const char*)pdata, plen, IGPU_OPT_POD_LIMITS, 0, &m) != IGPU_OK) {
    napi_throw_error(env, NULL, igpu_last_error(NULL));
    return NULL;
  }
  igpu_marshal_columns(m, &c);
  out = run_rollup(env, &c);
  if (out) { /* resource dictionary: id -> name, so the TypeScript side can rebuild Record<string,string> */
    napi_create_object(env, &names);
    for (i = 0; igpu_marshal_resource_name(m, i); ++i) {
      char key[16];
      snprintf(key, sizeof key, "%u", i);
      napi_create_string_utf8(env, igpu_marshal_resource_name(m, i), (size_t)-1, &v);
      napi_set_named_property(env, names, key, v);
    }
    napi_set_named_property(env, out, "resourceNames", names);
  }
  igpu_marshal_free(m);
  return out;
}

/* powerJoinSum({chipChip, chipInstance, energyChip, energyValue:Float64Array, powerMaxChip, powerMaxValue,
 *               unameInstance, unameNode, nChipIds, nInstanceIds}) */
static napi_value PowerJoinSum(napi_env env, napi_callback_info info) {
  size_t argc = 1, n = 0, tmp = 0;
  int bad = 0;
  napi_value arg, out, v;
  igpu_power_cols c;
  igpu_power_out o;
  CHECK(napi_get_cb_info(env, info, &argc, &arg, NULL, NULL));
  memset(&c, 0, sizeof c);
  memset(&o, 0, sizeof o);
  c.chip_chip = (const uint32_t*)ta_col(env, arg, "chipChip", napi_uint32_array, ANY_LEN, &n, &bad);
  c.n_chips = (uint32_t)n;
  c.chip_instance = (const uint32_t*)ta_col(env, arg, "chipInstance", napi_uint32_array, n, NULL, &bad);
  c.energy_chip = (const uint32_t*)ta_col(env, arg, "energyChip", napi_uint32_array, ANY_LEN, &tmp, &bad);
  c.n_energy = (uint32_t)tmp;
  c.energy_value = (const double*)ta_col(env, arg, "energyValue", napi_float64_array, tmp, NULL, &bad);
  c.power_max_chip = (const uint32_t*)ta_col(env, arg, "powerMaxChip", napi_uint32_array, ANY_LEN, &tmp, &bad);
  c.n_power_max = (uint32_t)tmp;
  c.power_max_value = (const double*)ta_col(env, arg, "powerMaxValue", napi_float64_array, tmp, NULL, &bad);
  c.uname_instance = (const uint32_t*)ta_col(env, arg, "unameInstance", napi_uint32_array, ANY_LEN, &tmp, &bad);
  c.n_uname = (uint32_t)tmp;
  c.uname_node = (const uint32_t*)ta_col(env, arg, "unameNode", napi_uint32_array, tmp, NULL, &bad);
  if (bad) return NULL;
  if (napi_get_named_property(env, arg, "nChipIds", &v) == napi_ok) napi_get_value_uint32(env, v, &c.n_chip_ids);
  if (napi_get_named_property(env, arg, "nInstanceIds", &v) == napi_ok) napi_get_value_uint32(env, v, &c.n_instance_ids);
  CHECK(napi_create_object(env, &out));
  new_ta(env, out, "chipFlags", napi_uint8_array, n, 1, (void**)&o.chip_flags);
  new_ta(env, out, "chipWatts", napi_float64_array, n, 8, (void**)&o.chip_watts);
  new_ta(env, out, "chipMaxWatts", napi_float64_array, n, 8, (void**)&o.chip_max_watts);
  new_ta(env, out, "chipNode", napi_uint32_array, n, 4, (void**)&o.chip_node);
  new_ta(env, out, "fleet", napi_float64_array, 4, 8, (void**)&o.fleet);
  if (igpu_power_join_sum(g_ctx, &c, &o) != IGPU_OK) {
    napi_throw_error(env, NULL, igpu_last_error(g_ctx));
    return NULL;
  }
  return out;
}

/* rangeView({x:Float32Array [series][nTime] (NaN = no sample), nodeOff:Uint32Array [nodes+1] first series of each node,
 *            nTime, window, q, fleetViews: 0|1})
 * -> {nodeQuantile, nodeSum, nodeCount:Float64Array [nodes]; nodeMin, nodeMax:Float32Array [nodes]; fleet:Float64Array [8]
 *     (sum, count, min, max, nodes reporting); with fleetViews: fleetQuantile:Float64Array [1], curveMin / curveMean /
 *     curveMax:Float32Array [windows], curveCount:Float64Array [windows]} */
static napi_value RangeView(napi_env env, napi_callback_info info) {
  size_t argc = 1, n_x = 0, n_off = 0;
  int bad = 0;
  napi_value arg, out = NULL, v;
  uint32_t n_time = 0, window = 60, fleet_views = 0, i;
  double q = 0.95;
  igpu_series* s = NULL;
  igpu_node_series_stats* st = NULL;
  uint64_t* ccount = NULL;
  const char* err = NULL;
  CHECK(napi_get_cb_info(env, info, &argc, &arg, NULL, NULL));
  const float* x = (const float*)ta_col(env, arg, "x", napi_float32_array, ANY_LEN, &n_x, &bad);
  const uint32_t* off = (const uint32_t*)ta_col(env, arg, "nodeOff", napi_uint32_array, ANY_LEN, &n_off, &bad);
  if (bad) return NULL;
  if (napi_get_named_property(env, arg, "nTime", &v) == napi_ok) napi_get_value_uint32(env, v, &n_time);
  if (napi_get_named_property(env, arg, "window", &v) == napi_ok) napi_get_value_uint32(env, v, &window);
  if (napi_get_named_property(env, arg, "q", &v) == napi_ok) napi_get_value_double(env, v, &q);
  if (napi_get_named_property(env, arg, "fleetViews", &v) == napi_ok) napi_get_value_uint32(env, v, &fleet_views);
  if (!n_time || !n_off || n_x % n_time || off[n_off - 1] != n_x / n_time) {
    napi_throw_error(env, "ERR_INVALID_ARG_VALUE", "igpu addon: 'x' must hold nodeOff[nodes] series of nTime samples");
    return NULL;
  }
  const uint32_t n_series = (uint32_t)(n_x / n_time), n_nodes = (uint32_t)n_off - 1;
  const uint32_t n_win = window ? (n_time + window - 1) / window : 0;
  double *nq, *nsum, *ncount, *fleet, *fq, *cc;
  float *nmin, *nmax, *cmin, *cmean, *cmax;
  CHECK(napi_create_object(env, &out));
  new_ta(env, out, "nodeQuantile", napi_float64_array, n_nodes, 8, (void**)&nq);
  new_ta(env, out, "nodeSum", napi_float64_array, n_nodes, 8, (void**)&nsum);
  new_ta(env, out, "nodeCount", napi_float64_array, n_nodes, 8, (void**)&ncount);
  new_ta(env, out, "nodeMin", napi_float32_array, n_nodes, 4, (void**)&nmin);
  new_ta(env, out, "nodeMax", napi_float32_array, n_nodes, 4, (void**)&nmax);
  new_ta(env, out, "fleet", napi_float64_array, IGPU_FLEET_LEN, 8, (void**)&fleet);
  st = (igpu_node_series_stats*)malloc((n_nodes ? n_nodes : 1) * sizeof *st);
  if (!st) err = "igpu addon: out of memory";
  if (!err && igpu_series_create(g_ctx, n_series, n_time, n_nodes, off, &s) != IGPU_OK) err = igpu_last_error(g_ctx);
  if (!err && igpu_series_upload(g_ctx, s, 0, n_series, x) != IGPU_OK) err = igpu_last_error(g_ctx);
  if (!err && fleet_views && igpu_series_keep_window_counts(g_ctx, s, 1) != IGPU_OK) err = igpu_last_error(g_ctx);
  /* window planes stay on the device (the page draws the fleet curve, not 1,440 points per chip) */
  if (!err && igpu_series_rollup(g_ctx, s, window, q, 0, NULL, NULL, NULL, nq, st, fleet) != IGPU_OK) err = igpu_last_error(g_ctx);
  for (i = 0; !err && i < n_nodes; ++i) nsum[i] = st[i].sum, ncount[i] = (double)st[i].count, nmin[i] = st[i].min, nmax[i] = st[i].max;
  if (!err && fleet_views) {
    igpu_window_rollup_out wo;
    memset(&wo, 0, sizeof wo);
    new_ta(env, out, "fleetQuantile", napi_float64_array, 1, 8, (void**)&fq);
    new_ta(env, out, "curveMin", napi_float32_array, n_win, 4, (void**)&cmin);
    new_ta(env, out, "curveMean", napi_float32_array, n_win, 4, (void**)&cmean);
    new_ta(env, out, "curveMax", napi_float32_array, n_win, 4, (void**)&cmax);
    new_ta(env, out, "curveCount", napi_float64_array, n_win, 8, (void**)&cc);
    ccount = (uint64_t*)malloc((n_win ? n_win : 1) * sizeof *ccount);
    wo.fleet_min = cmin, wo.fleet_mean = cmean, wo.fleet_max = cmax, wo.fleet_count = ccount;
    if (!ccount) err = "igpu addon: out of memory";
    if (!err && igpu_series_quantile_fleet(g_ctx, s, q, 0, fq, NULL) != IGPU_OK) err = igpu_last_error(g_ctx);
    if (!err && igpu_series_window_rollup(g_ctx, s, window, &wo) != IGPU_OK) err = igpu_last_error(g_ctx);
    for (i = 0; !err && i < n_win; ++i) cc[i] = (double)ccount[i];
  }
  if (err) napi_throw_error(env, NULL, err); /* (the message is copied before the handle that owns it goes away) */
  if (s) igpu_series_destroy(g_ctx, s);
  free(st);
  free(ccount);
  return err ? NULL : out;
}

NAPI_MODULE_INIT() {
  napi_property_descriptor props[] = {
      {"rollupFleet", NULL, RollupFleet, NULL, NULL, NULL, 0, NULL},
      {"rollupFleetJson", NULL, RollupFleetJson, NULL, NULL, NULL, 0, NULL},
      {"powerJoinSum", NULL, PowerJoinSum, NULL, NULL, NULL, 0, NULL},
      {"rangeView", NULL, RangeView, NULL, NULL, NULL, 0, NULL},
  };
  if (igpu_ctx_create(0, NULL, &g_ctx) != IGPU_OK) {
    napi_throw_error(env, NULL, igpu_last_error(NULL)); /* no B200: the TypeScript layer stays on the reference path */
    return NULL;
  }
  if (napi_define_properties(env, exports, sizeof props / sizeof props[0], props) != napi_ok) return NULL;
  return exports;
}
