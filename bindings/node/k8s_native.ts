/**
 * k8s_native.ts — Node-side twin of headlamp_intel_gpu_plugin_b200/k8s.py (SURVEY.md §8(f) N2).
 *
 * UNVERIFIED: there is no Node.js / tsc in the build image, so this file has never been compiled or run.
 * It shows how src/api/k8s.ts keeps its exported signatures (SURVEY §8b) while the arithmetic moves to
 * libigpu_rollup.so through the N-API addon (igpu_addon.c).  In a browser bundle `native` is undefined and
 * every function below falls through to the reference implementation, which is all a browser can do.
 *
 * The columnar decode mirrors FleetRollup in k8s.py: planes are column-major (`podReq[r * nPods + p]`),
 * masks say which keys exist, integers come back as BigInt64Array (exact; the reference's float64 sums are
 * exact below 2^53, which the marshal layer enforces).
 */
import * as ref from './k8s'; // the reference module, unchanged

type Addon = {
  rollupFleetJson(nodesJson: Uint8Array, podsJson: Uint8Array): RollupOut;
};
type RollupOut = {
  podIsGpu: Uint8Array; podReqMask: Uint32Array; podReq: BigInt64Array;
  nodeOutFlags: Uint8Array; nodeType: Uint8Array; nodeGpuCount: BigInt64Array;
  nodeCapTotal: BigInt64Array; nodeAllocTotal: BigInt64Array; nodePodCount: Uint32Array;
  nodeBarPct: BigInt64Array; nodeBarLevel: Uint8Array; nodeInUse: BigInt64Array;
  nodeUtilPct: BigInt64Array; nodeUtilLevel: Uint8Array; nodeReq: BigInt64Array;
  fleetTotals: BigInt64Array; resourceNames: Record<string, string>;
};

let native: Addon | undefined;
try {
  // eslint-disable-next-line @typescript-eslint/no-var-requires
  native = typeof process !== 'undefined' && process.versions?.node ? require('./igpu_addon.node') : undefined;
} catch {
  native = undefined; // no B200 / addon not built: stay on the reference path
}

const enc = new TextEncoder();
const GPU_TYPES: ref.GpuType[] = ['discrete', 'integrated', 'unknown'];
const NODEOUT_IS_GPU = 0x01, NODEOUT_READY = 0x02, NODEOUT_CAP_NAN = 0x04, NODEOUT_ALLOC_NAN = 0x08, NODEOUT_BAR_DASH = 0x10;
const T = { gpuNodes: 0, discrete: 1, integrated: 2, unknown: 3, gpuDevices: 4, readyNodes: 5, capacity: 6, allocatable: 7,
            capNaN: 8, allocNaN: 9, allocated: 10, gpuPods: 11, phase0: 12, utilPct: 18 } as const;

/** The added batch entry point: one GPU rollup for the whole fleet (IntelGpuDataContext would call it once per refresh). */
export function rollupFleet(nodes: unknown[], pods: unknown[]) {
  if (!native) return undefined;
  const out = native.rollupFleetJson(enc.encode(JSON.stringify(nodes)), enc.encode(JSON.stringify(pods)));
  const R = Object.keys(out.resourceNames).length, P = pods.length;
  const t = out.fleetTotals, n = (i: number) => Number(t[i]);
  return {
    out,
    gpuNodes: nodes.filter((_, i) => out.nodeOutFlags[i] & NODEOUT_IS_GPU) as ref.IntelGpuNode[],
    gpuPods: pods.filter((_, p) => out.podIsGpu[p]) as ref.IntelGpuPod[],
    podRequests(p: number): Record<string, string> { // getPodGpuRequests, k8s.ts:289-301
      const res: Record<string, string> = {};
      for (let r = 0; r < R; r++) if ((out.podReqMask[p] >>> r) & 1) res[out.resourceNames[r]] = String(out.podReq[r * P + p]);
      return res;
    },
    overview: { // OverviewPage.tsx:71-130
      discreteCount: n(T.discrete), integratedCount: n(T.integrated), unknownCount: n(T.unknown),
      totalGpuCount: n(T.gpuDevices), readyNodeCount: n(T.readyNodes), totalGpuNodes: n(T.gpuNodes),
      totalCapacityGpus: n(T.capNaN) ? NaN : n(T.capacity), totalAllocatableGpus: n(T.allocNaN) ? NaN : n(T.allocatable),
      totalAllocatedGpus: n(T.allocated), gpuUtilizationPct: n(T.utilPct),
      podPhaseCounts: { Running: n(T.phase0), Pending: n(T.phase0 + 1), Succeeded: n(T.phase0 + 2), Failed: n(T.phase0 + 3), Other: n(T.phase0 + 4) },
    },
    nodeRow(i: number) { // NodesPage.tsx:162-189, :35-38 and NodeDetailSection.tsx:69-91
      const f = out.nodeOutFlags[i];
      return {
        gpuType: GPU_TYPES[out.nodeType[i]], gpuCount: Number(out.nodeGpuCount[i]), ready: !!(f & NODEOUT_READY),
        totalCapacity: f & NODEOUT_CAP_NAN ? NaN : Number(out.nodeCapTotal[i]),
        totalAllocatable: f & NODEOUT_ALLOC_NAN ? NaN : Number(out.nodeAllocTotal[i]),
        podCount: out.nodePodCount[i],
        bar: f & NODEOUT_BAR_DASH ? null : { pct: Number(out.nodeBarPct[i]), level: out.nodeBarLevel[i] },
        gpuInUse: Number(out.nodeInUse[i]), utilizationPct: Number(out.nodeUtilPct[i]), utilizationLevel: out.nodeUtilLevel[i],
      };
    },
  };
}

// ---- frozen signatures: same names, parameters and return types as src/api/k8s.ts ------------------------
export function isIntelGpuNode(node: unknown): node is ref.IntelGpuNode {
  if (!native || !node || typeof node !== 'object') return ref.isIntelGpuNode(node);
  return !!(rollupFleet([node], [])!.out.nodeOutFlags[0] & NODEOUT_IS_GPU);
}
export function filterIntelGpuNodes(items: unknown[]): ref.IntelGpuNode[] {
  return native && items.length ? rollupFleet(items, [])!.gpuNodes : ref.filterIntelGpuNodes(items);
}
export function getNodeGpuCount(node: ref.IntelGpuNode): number {
  return native ? Number(rollupFleet([node], [])!.out.nodeGpuCount[0]) : ref.getNodeGpuCount(node);
}
export function getNodeGpuType(node: ref.IntelGpuNode): ref.GpuType {
  return native ? GPU_TYPES[rollupFleet([node], [])!.out.nodeType[0]] : ref.getNodeGpuType(node);
}
export function isNodeReady(node: ref.IntelGpuNode): boolean {
  return native ? !!(rollupFleet([node], [])!.out.nodeOutFlags[0] & NODEOUT_READY) : ref.isNodeReady(node);
}
export function isGpuRequestingPod(pod: unknown): pod is ref.IntelGpuPod {
  if (!native || !pod || typeof pod !== 'object') return ref.isGpuRequestingPod(pod);
  return !!rollupFleet([], [pod])!.out.podIsGpu[0];
}
export function filterGpuRequestingPods(items: unknown[]): ref.IntelGpuPod[] {
  return native && items.length ? rollupFleet([], items)!.gpuPods : ref.filterGpuRequestingPods(items);
}
export function getPodGpuRequests(pod: ref.IntelGpuPod): Record<string, string> {
  return native ? rollupFleet([], [pod])!.podRequests(0) : ref.getPodGpuRequests(pod);
}
// presentation helpers and constants are re-exported untouched
export { formatAge, formatGpuResourceName, formatGpuType, getGpuResources, isKubeList, pluginStatusText, pluginStatusToStatus } from './k8s';
