// CudaBackend for TStwo over bun:ffi -- the reference-side binding of include/tstwo_cuda.h.
// NOT executable in the build image (no bun); it documents exactly what a TStwo maintainer adds under
// packages/core/src/backend/cuda/.  The Python mirror tstwo_b200/backend.py has the same classes and is what the
// parity tests exercise.  Names follow packages/core/src/backend/cpu/{index,circle,blake2,fri}.ts.
import { dlopen, FFIType, ptr, suffix } from "bun:ffi";
import { CircleEvaluation, BitReversedOrder } from "../../poly/circle/evaluation";
import { CirclePoly } from "../../poly/circle/poly";
import type { CircleDomain } from "../../poly/circle/domain";
import { CanonicCoset } from "../../poly/circle/canonic";
import type { Coset } from "../../circle";

const T = FFIType;
const lib = dlopen(`libtstwo_cuda.${suffix}`, {
  tstwo_init: { args: [T.i32], returns: T.i32 },
  tstwo_last_error: { args: [], returns: T.cstring },
  tstwo_malloc: { args: [T.u64, T.ptr], returns: T.i32 },
  tstwo_free: { args: [T.u64], returns: T.i32 },
  tstwo_upload: { args: [T.u64, T.ptr, T.u64], returns: T.i32 },
  tstwo_download: { args: [T.ptr, T.u64, T.u64], returns: T.i32 },
  tstwo_bit_reverse: { args: [T.u64, T.u64], returns: T.i32 },
  tstwo_twiddles_precompute: { args: [T.u32, T.i32, T.ptr], returns: T.i32 },
  tstwo_interpolate_columns: { args: [T.ptr, T.ptr, T.i32, T.i32, T.u32, T.u64, T.u32], returns: T.i32 },
  tstwo_evaluate_columns: { args: [T.ptr, T.i32, T.ptr, T.i32, T.i32, T.u32, T.u64, T.u32], returns: T.i32 },
  tstwo_merkle_commit_layer: { args: [T.i32, T.u64, T.ptr, T.i32, T.u64], returns: T.i32 },
  tstwo_merkle_commit: { args: [T.ptr, T.ptr, T.i32, T.u64, T.ptr], returns: T.i32 },
  tstwo_merkle_leaf_absorb: { args: [T.u64, T.u64, T.i32, T.i32, T.i32, T.i32, T.u64], returns: T.i32 },
  tstwo_merkle_commit_nodes: { args: [T.i32, T.u64, T.ptr], returns: T.i32 },
  tstwo_fold_line: { args: [T.ptr, T.i32, T.u32, T.ptr, T.ptr, T.u64], returns: T.i32 },
  tstwo_fold_circle_into_line: { args: [T.ptr, T.ptr, T.i32, T.u32, T.ptr, T.u64], returns: T.i32 },
  tstwo_lde_commit: { args: [T.u64, T.i32, T.i32, T.i32, T.u64, T.u64, T.u64, T.u64, T.ptr, T.u32], returns: T.i32 },
  tstwo_lde_commit_host: { args: [T.ptr, T.i32, T.i32, T.i32, T.u64, T.u64, T.u64, T.u64, T.u64, T.ptr, T.u32], returns: T.i32 },
  // multi-GPU, one process per GPU (the torchrun path of bench.py): last FFT pass stores into the peers' leaf buffers, mapped
  // through CUDA IPC.  A single Bun process uses tstwo_init_devices + tstwo_lde_commit_multi below instead.
  tstwo_lde_columns_to_peers: { args: [T.u64, T.i32, T.i32, T.i32, T.u64, T.u64, T.u64, T.ptr, T.i32, T.i32, T.u32], returns: T.i32 },
  tstwo_lde_columns_to_peers_overlapped: { args: [T.u64, T.i32, T.i32, T.i32, T.u64, T.u64, T.u64, T.ptr, T.i32, T.i32, T.u32, T.i32, T.i32], returns: T.i32 },
  tstwo_lde_columns_to_peers_halves: { args: [T.u64, T.i32, T.i32, T.i32, T.u64, T.u64, T.u64, T.ptr, T.i32, T.i32, T.i32, T.u32, T.i32, T.i32, T.ptr, T.ptr, T.i32, T.u32], returns: T.i32 },
  tstwo_peer_join: { args: [], returns: T.i32 },
  tstwo_ipc_export: { args: [T.u64, T.ptr], returns: T.i32 },
  tstwo_ipc_import: { args: [T.ptr, T.ptr], returns: T.i32 },
  tstwo_ipc_close: { args: [T.u64], returns: T.i32 },
  tstwo_peer_signal: { args: [T.ptr, T.i32, T.u32], returns: T.i32 },
  tstwo_peer_wait: { args: [T.u64, T.i32, T.u32, T.u32], returns: T.i32 },
  tstwo_peer_put: { args: [T.u64, T.u64, T.i32, T.u64, T.u32], returns: T.i32 },
  // decommitment witnesses, proof of work, Stwo-correct DEEP quotients (callers: MerkleProver.decommit, FriProver.decommit, prove_values)
  tstwo_gather_words: { args: [T.ptr, T.i32, T.i32, T.ptr], returns: T.i32 },
  tstwo_grind: { args: [T.ptr, T.u32, T.ptr], returns: T.i32 },
  tstwo_merkle_decommit: { args: [T.u64, T.i32, T.ptr, T.ptr, T.i32, T.ptr, T.ptr, T.ptr, T.ptr, T.ptr, T.ptr], returns: T.i32 },
  tstwo_fri_decommit_layers: { args: [T.i32, T.ptr, T.ptr, T.ptr, T.ptr, T.i32, T.ptr, T.u64, T.ptr, T.u64, T.ptr, T.u64, T.ptr], returns: T.i32 },
  tstwo_accumulate_quotients_ex: { args: [T.i32, T.u32, T.ptr, T.i32, T.ptr, T.i32, T.ptr, T.ptr, T.ptr, T.ptr, T.ptr, T.u32], returns: T.i32 },
  // wide-Fibonacci constraint quotients over the evaluation domain (ComponentProver.evaluateConstraintQuotientsOnDomain); allocator trim
  tstwo_widefib_constraints: { args: [T.i32, T.i32, T.ptr, T.i32, T.ptr, T.ptr, T.ptr], returns: T.i32 },
  tstwo_mem_trim: { args: [], returns: T.i32 },
  // the rest of include/tstwo_cuda.h (tests/test_abi_cpu.py keeps this table and the header in step)
  tstwo_init_devices: { args: [T.i32, T.ptr], returns: T.i32 },
  tstwo_device_count: { args: [], returns: T.i32 },
  tstwo_set_device: { args: [T.i32], returns: T.i32 },
  tstwo_shutdown: { args: [], returns: T.i32 },
  tstwo_set_stream: { args: [T.ptr], returns: T.i32 },
  tstwo_sync: { args: [], returns: T.i32 },
  tstwo_device_info: { args: [T.ptr, T.ptr, T.ptr, T.ptr], returns: T.i32 },
  tstwo_launch_count: { args: [], returns: T.u64 },
  tstwo_memset_zero: { args: [T.u64, T.u64], returns: T.i32 },
  tstwo_copy: { args: [T.u64, T.u64, T.u64], returns: T.i32 },
  tstwo_twiddles_free: { args: [T.u64], returns: T.i32 },
  tstwo_twiddles_info: { args: [T.u64, T.ptr, T.ptr, T.ptr, T.ptr], returns: T.i32 },
  tstwo_eval_at_point: { args: [T.u64, T.i32, T.ptr, T.ptr, T.ptr], returns: T.i32 },
  tstwo_eval_at_point_many: { args: [T.ptr, T.i32, T.i32, T.ptr, T.ptr, T.ptr], returns: T.i32 },
  tstwo_eval_at_point_noalloc: { args: [T.u64, T.i32, T.ptr, T.ptr, T.ptr], returns: T.i32 },
  tstwo_eval_at_point_noalloc_max_log: { args: [], returns: T.i32 },
  tstwo_fri_decompose: { args: [T.ptr, T.i32, T.ptr, T.ptr], returns: T.i32 },
  tstwo_accumulate: { args: [T.ptr, T.ptr, T.u64], returns: T.i32 },
  tstwo_accumulate_quotients: { args: [T.i32, T.u32, T.ptr, T.i32, T.ptr, T.i32, T.ptr, T.ptr, T.ptr, T.ptr, T.ptr], returns: T.i32 },
  tstwo_lde_columns_sharded: { args: [T.u64, T.i32, T.i32, T.i32, T.u64, T.u64, T.u64, T.i32, T.u32], returns: T.i32 },
  tstwo_scatter_shards: { args: [T.u64, T.i32, T.u64, T.ptr, T.i32, T.i32], returns: T.i32 },
  tstwo_scatter_join: { args: [], returns: T.i32 },
  tstwo_fri_commit_layers: { args: [T.ptr, T.i32, T.u32, T.i32, T.ptr, T.ptr, T.ptr, T.i32, T.u64, T.ptr, T.ptr, T.ptr, T.ptr, T.ptr, T.ptr], returns: T.i32 },
  tstwo_lde_commit_multi: { args: [T.ptr, T.i32, T.i32, T.i32, T.ptr, T.u32], returns: T.i32 },
  tstwo_lde_commit_multi_dev: { args: [T.ptr, T.i32, T.i32, T.i32, T.ptr, T.u32], returns: T.i32 },
}).symbols;

function check(rc: number): void {
  if (rc !== 0) throw new Error(String(lib.tstwo_last_error()));   // same messages as the CpuBackend throws
}
check(lib.tstwo_init(Number(process.env.TSTWO_CUDA_DEVICE ?? 0)));
const QUIRKS = process.env.TSTWO_QUIRKS === "0" ? 0 : 1;            // log-3 swap of backend/cpu/circle.ts:127-131

/** Column<BaseField> resident on the device (backend/index.ts:53-74). */
export class CudaColumn {
  constructor(readonly dptr: bigint, readonly length: number) {}
  static fromArray(values: Uint32Array): CudaColumn {
    const out = new BigUint64Array(1);
    check(lib.tstwo_malloc(BigInt(values.byteLength), ptr(out)));
    check(lib.tstwo_upload(out[0], ptr(values), BigInt(values.byteLength)));
    return new CudaColumn(out[0], values.length);
  }
  len(): number { return this.length; }
  toCpu(): Uint32Array {
    const host = new Uint32Array(this.length);
    check(lib.tstwo_download(ptr(host), this.dptr, BigInt(host.byteLength)));
    return host;
  }
  at(i: number): number {
    const one = new Uint32Array(1);
    check(lib.tstwo_download(ptr(one), this.dptr + BigInt(4 * i), 4n));
    return one[0];
  }
}

export class CudaTwiddleTree { constructor(readonly rootCoset: Coset, readonly handle: bigint) {} }

function addrs(cols: CudaColumn[]): BigUint64Array { return BigUint64Array.from(cols.map((c) => c.dptr)); }

/** Same statics as CpuCircleEvaluation (backend/cpu/circle.ts:17-37): duck-typed dispatch of evaluation.ts:138-166. */
// @ts-expect-error -- as in the reference's CpuCircleEvaluation
export class CudaCircleEvaluation extends CircleEvaluation<any, any, BitReversedOrder> {
  static precomputeTwiddles(coset: Coset): CudaTwiddleTree {
    const h = new BigUint64Array(1);
    check(lib.tstwo_twiddles_precompute(coset.initial_index.value, coset.log_size, ptr(h)));
    return new CudaTwiddleTree(coset, h[0]);
  }
  static bitReverseColumn(col: CudaColumn): void { check(lib.tstwo_bit_reverse(col.dptr, BigInt(col.length))); }
  static to_cpu(values: CudaColumn): Uint32Array { return values.toCpu(); }
  /** CpuCirclePoly.interpolate (backend/cpu/circle.ts:136-207), batched over columns of one domain. */
  static interpolateColumns(evals: CudaCircleEvaluation[], tw: CudaTwiddleTree): CudaCirclePoly[] {
    const d = evals[0].domain, n = d.size();
    const out = evals.map(() => CudaColumn.fromArray(new Uint32Array(n)));
    check(lib.tstwo_interpolate_columns(ptr(addrs(evals.map((e) => e.values as unknown as CudaColumn))), ptr(addrs(out)),
      evals.length, d.log_size(), d.halfCoset.initial_index.value, tw.handle, QUIRKS));
    return out.map((c) => new CudaCirclePoly(c as any));
  }
}

export class CudaCirclePoly extends CirclePoly<any> {
  static precomputeTwiddles = CudaCircleEvaluation.precomputeTwiddles;
  /** CpuCirclePoly.evaluate with extend (backend/cpu/circle.ts:71-134). */
  static evaluate(poly: CudaCirclePoly, domain: CircleDomain, tw: CudaTwiddleTree): CudaCircleEvaluation {
    const src = poly.coeffs as unknown as CudaColumn;
    const dst = CudaColumn.fromArray(new Uint32Array(domain.size()));
    check(lib.tstwo_evaluate_columns(ptr(addrs([src])), Math.log2(src.length), ptr(addrs([dst])), 1, domain.log_size(),
      domain.halfCoset.initial_index.value, tw.handle, QUIRKS));
    return new CudaCircleEvaluation(domain, dst as any);
  }
}

/** MerkleOps<Blake2sHash> (vcs/ops.ts:16-26): hashes stay on the device as a CudaColumn of 8 words per hash. */
export class CudaMerkleOps {
  commitOnLayer(logSize: number, prevLayer: CudaColumn | undefined, columns: CudaColumn[]): CudaColumn {
    const out = CudaColumn.fromArray(new Uint32Array(8 << logSize));
    check(lib.tstwo_merkle_commit_layer(logSize, prevLayer ? prevLayer.dptr : 0n, ptr(addrs(columns)), columns.length, out.dptr));
    return out;
  }
}

/** Stwo's CommitmentTreeProver::new (Rust text at pcs/prover.ts:215-235), fused on the device. */
export function commitTrace(trace: CudaColumn, nCols: number, logSize: number, logBlowup: number, tw: CudaTwiddleTree) {
  const n = 1 << logSize, ne = n << logBlowup, ext = logSize + logBlowup;
  const coeffs = CudaColumn.fromArray(new Uint32Array(0)), root = new Uint8Array(32);
  const evals = new BigUint64Array(1), layers = new BigUint64Array(1), cf = new BigUint64Array(1);
  check(lib.tstwo_malloc(BigInt(4 * nCols * n), ptr(cf)));
  check(lib.tstwo_malloc(BigInt(4 * nCols * ne), ptr(evals)));
  check(lib.tstwo_malloc(BigInt(32 * ((2 << ext) - 1)), ptr(layers)));
  check(lib.tstwo_lde_commit(trace.dptr, nCols, logSize, logBlowup, tw.handle, cf[0], evals[0], layers[0], ptr(root), QUIRKS));
  void coeffs;
  return { root, coeffs: cf[0], evals: evals[0], layers: layers[0], domain: CanonicCoset.new(ext).circleDomain() };
}

/** The same commitment over ALL GPUs of the box from this one Bun thread (BASELINE config 3): `devices` = CUDA ordinals, a power
 * of two dividing nCols.  The trace is a host Uint32Array [nCols][2^logSize]; the library shards it by column, pulls every shard
 * over its own PCIe link, exchanges the evaluations over NVLink peer memory inside the last FFT pass, builds one Merkle subtree
 * per device and finishes the top layers on device 0.  The root equals commitTrace()'s bit for bit. */
let multiDevices = 0;
export function commitTraceMulti(trace: Uint32Array, nCols: number, logSize: number, logBlowup: number, devices: number[]): Uint8Array {
  if (multiDevices !== devices.length) {
    check(lib.tstwo_init_devices(devices.length, ptr(Int32Array.from(devices))));   // slot 0 stays current: single-device calls keep working
    multiDevices = devices.length;
  }
  const root = new Uint8Array(32);
  check(lib.tstwo_lde_commit_multi(ptr(trace), nCols, logSize, logBlowup, ptr(root), QUIRKS));
  return root;
}

// GrindOps for the CUDA backend (backend/cpu/grind.ts:31-42): the nonce search runs on the device, the channel stays on the host.
export function grind(channel: { digest(): { bytes: Uint8Array } }, powBits: number): bigint {
  const nonce = new BigUint64Array(1);
  check(lib.tstwo_grind(ptr(channel.digest().bytes), powBits, ptr(nonce)));
  return nonce[0];
}

// The witness half of MerkleProver.decommit / computeDecommitmentPositionsAndWitnessEvals: the index walk stays in TS
// (vcs/prover.ts:32-109 depends on the queries only), the words come from device memory in one launch.
export function gatherWords(addresses: bigint[], wordsEach: number): Uint32Array {
  const out = new Uint32Array(addresses.length * wordsEach);
  if (addresses.length) check(lib.tstwo_gather_words(ptr(BigUint64Array.from(addresses)), addresses.length, wordsEach, ptr(out)));
  return out;
}

// ComponentProver.evaluateConstraintQuotientsOnDomain for the wide-Fibonacci component (the loop that
// constraint_framework/index.ts:131-152 leaves as a placeholder): `evals` are the trace columns on
// CanonicCoset(logSize + 1).circleDomain(); `coeffs` = accum.randomCoeffPowers reversed (constraint i gets alpha^(N-1-i),
// air/accumulator.ts:159-176); `denomInv[i]` = 1 / cosetVanishing(CanonicCoset(logSize).coset, evalDomain.at(i)), bit-reversed.
// Adds denomInv[row >> logSize] * sum_i coeffs[i] * (c - a^2 - b^2)_i(row) into the accumulator's 4 coordinate columns.
export function wideFibonacciConstraintQuotients(logSize: number, evals: CudaColumn[], coeffs: Uint32Array, denomInv: Uint32Array,
                                                 accCoords: [CudaColumn, CudaColumn, CudaColumn, CudaColumn]): void {
  const cols = BigUint64Array.from(evals.map((c) => c.dptr));
  const out = BigUint64Array.from(accCoords.map((c) => c.dptr));
  check(lib.tstwo_widefib_constraints(logSize, logSize + 1, ptr(cols), evals.length, ptr(coeffs), ptr(denomInv), ptr(out)));
}

// MerkleProver.decommit (vcs/prover.ts:32-109) for a tree written by tstwo_merkle_commit: queriesPerLogSize as in the reference.
export function merkleDecommit(tree: bigint, maxLog: number, columns: CudaColumn[], queriesPerLogSize: Map<number, number[]>):
    { queried: Uint32Array; hashWitness: Uint8Array[]; columnWitness: Uint32Array } {
  const offsets = new Int32Array(maxLog + 2), flat: number[] = [];
  for (let log = 0; log <= maxLog; log++) { flat.push(...(queriesPerLogSize.get(log) ?? [])); offsets[log + 1] = flat.length; }
  const cols = BigUint64Array.from(columns.map((c) => c.dptr)), logs = Int32Array.from(columns.map((c) => Math.log2(c.length)));
  const positions = Uint32Array.from(flat.length ? flat : [0]), counts = new BigUint64Array(3);
  const call = (q: Uint32Array | null, h: Uint8Array | null, w: Uint32Array | null) =>
    check(lib.tstwo_merkle_decommit(tree, maxLog, ptr(cols), ptr(logs), columns.length, ptr(offsets), ptr(positions),
                                    q ? ptr(q) : null, h ? ptr(h) : null, w ? ptr(w) : null, ptr(counts)));
  call(null, null, null);                                          // sizing pass
  const queried = new Uint32Array(Math.max(1, Number(counts[0]))), hashes = new Uint8Array(Math.max(1, 32 * Number(counts[1])));
  const witness = new Uint32Array(Math.max(1, Number(counts[2])));
  call(queried, hashes, witness);
  const hashWitness = Array.from({ length: Number(counts[1]) }, (_, i) => hashes.slice(32 * i, 32 * i + 32));
  return { queried: queried.subarray(0, Number(counts[0])), hashWitness, columnWitness: witness.subarray(0, Number(counts[2])) };
}

// FriProver.commitInnerLayers (fri.ts:676-716) in one call with the Fiat-Shamir transcript on the device.  `state` is the
// Blake2sChannel's state as 19 words (digest[8], n_challenges, n_sent, queue length, queued base felts[8]); it is updated in
// place and the caller writes it back into its channel object.  Returns per layer the tree buffer, root, folding alpha and the
// next layer's four coordinate columns (quarters of ONE device block: free evals[l][0] only).
export function friCommitInnerLayers(first: [CudaColumn, CudaColumn, CudaColumn, CudaColumn], firstLog: number, firstCosetInit: number,
                                     lastLog: number, circle: { coords: CudaColumn[]; log: number; halfInit: number }[], tw: bigint,
                                     state: Uint32Array) {
  const n = Math.max(0, firstLog - lastLog);
  const f = BigUint64Array.from(first.map((c) => c.dptr));
  const cc = BigUint64Array.from(circle.flatMap((c) => c.coords.map((k) => k.dptr)));
  const clogs = Int32Array.from(circle.map((c) => c.log)), cinits = Uint32Array.from(circle.map((c) => c.halfInit));
  const evals = new BigUint64Array(4 * Math.max(1, n)), trees = new BigUint64Array(Math.max(1, n));
  const roots = new Uint8Array(32 * Math.max(1, n)), alphas = new Uint32Array(4 * Math.max(1, n)), nOut = new Int32Array(1);
  check(lib.tstwo_fri_commit_layers(ptr(f), firstLog, firstCosetInit, lastLog, circle.length ? ptr(cc) : null, circle.length ? ptr(clogs) : null,
                                    circle.length ? ptr(cinits) : null, circle.length, tw, ptr(state), ptr(evals), ptr(trees), ptr(roots),
                                    ptr(alphas), ptr(nOut)));
  return Array.from({ length: nOut[0] }, (_, l) => ({
    tree: trees[l], root: roots.slice(32 * l, 32 * l + 32), alpha: alphas.slice(4 * l, 4 * l + 4), nextEvals: Array.from(evals.slice(4 * l, 4 * l + 4)) }));
}

// FriProver.decommitOnQueries' loop over the inner layers (fri.ts:769-785) in one call: `queries` are the sorted positions on the
// first inner layer's domain.  Returns per layer { friWitness (QM31 as 4 words each), hashWitness, columnWitness }.
export function friDecommitInnerLayers(layers: { tree: bigint; coords: CudaColumn[]; log: number }[], queries: number[]) {
  const n = layers.length, nq = queries.length;
  const trees = BigUint64Array.from(layers.map((l) => l.tree)), evals = BigUint64Array.from(layers.flatMap((l) => l.coords.map((c) => c.dptr)));
  const logs = Int32Array.from(layers.map((l) => l.log)), q = Uint32Array.from(nq ? queries : [0]);
  const capW = Math.max(1, nq * n), capH = Math.max(1, 2 * nq * layers.reduce((a, l) => a + l.log, 0)), capC = Math.max(1, 8 * nq * n);
  const wit = new Uint32Array(4 * capW), hashes = new Uint8Array(32 * capH), colw = new Uint32Array(capC), counts = new BigUint64Array(3 * Math.max(1, n));
  check(lib.tstwo_fri_decommit_layers(n, ptr(trees), ptr(evals), ptr(logs), ptr(q), nq, ptr(wit), capW, ptr(hashes), capH, ptr(colw), capC, ptr(counts)));
  let wi = 0, hi = 0, ci = 0;
  return layers.map((_, l) => {
    const nw = Number(counts[3 * l]), nh = Number(counts[3 * l + 1]), nc = Number(counts[3 * l + 2]);
    const out = { friWitness: wit.slice(4 * wi, 4 * (wi + nw)), hashWitness: Array.from({ length: nh }, (_, k) => hashes.slice(32 * (hi + k), 32 * (hi + k) + 32)),
                  columnWitness: colw.slice(ci, ci + nc) };
    wi += nw; hi += nh; ci += nc;
    return out;
  });
}
