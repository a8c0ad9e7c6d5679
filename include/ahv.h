/*
 * ahv.h - C ABI of libahv_b200.so: the B200-native replacement for the one
 * data-parallel hot path of imran-siddique/agent-hypervisor (Delta Audit
 * Engine: SHA-256 leaf hashing, Merkle roots, chain verification; plus the
 * sigma_eff + execution-ring gating pass).
 *
 * The reference has no FFI (it is pure Python); the interface each entry point
 * replaces is therefore a Python method, cited as file:line under
 * /root/reference/src/hypervisor/.  INTEGRATION.md shows the ctypes binding a
 * maintainer would add on the reference side.
 *
 * Conventions
 *   - every function returns 0 on success, <0 (AHV_E_*) on error;
 *     ahv_last_error() returns a thread-local message for the last failure.
 *   - the caller owns every buffer it passes.  The library owns only the
 *     device scratch attached to an ahv_ctx (one ctx per GPU per process).
 *   - "host" entry points take HOST pointers, copy in/out themselves and
 *     return after the results are in the caller's buffers.
 *   - "_dev" entry points take DEVICE pointers, enqueue on the ctx stream and
 *     return without synchronising (call ahv_sync).
 *   - digests are raw 32-byte SHA-256 values (hashlib .digest() order).
 *   - a packed arena is `uint8 arena[]` + `uint64 offsets[n+1]`; message i is
 *     arena[offsets[i] .. offsets[i+1]).  Sessions are CSR ranges of message
 *     indices: `uint64 session_rowptr[n_sessions+1]`.
 *   - no CPU fallback exists: without a CUDA device ahv_create fails.
 */
#ifndef AHV_H
#define AHV_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AHV_ABI_VERSION 2

#define AHV_OK 0
#define AHV_E_INVALID (-1)  /* bad argument */
#define AHV_E_CUDA (-2)     /* CUDA runtime error (message has the detail) */
#define AHV_E_NOMEM (-3)    /* device or pinned-host allocation failed */
#define AHV_E_NODEVICE (-4) /* no usable CUDA device */
#define AHV_E_TOOLARGE (-5) /* size beyond a documented limit */

/* parent_kind values (host encoding of SemanticDelta.parent_hash, audit/delta.py:38) */
#define AHV_PARENT_NONE 0  /* parent_hash is None */
#define AHV_PARENT_HEX 1   /* 64 lowercase hex chars; raw 32 bytes in parent_digests */
#define AHV_PARENT_OTHER 2 /* any other string: never equal to a hexdigest */

/* action_flags bits (host encoding of ActionDescriptor, models.py:104-132) */
#define AHV_ACT_ADMIN 1
#define AHV_ACT_READ_ONLY 2
#define AHV_ACT_REV_NONE 4

/* RingEnforcer.check outcome codes, in the order rings/enforcer.py:58-128 tests them */
#define AHV_REASON_GRANTED 0
#define AHV_REASON_NEEDS_WITNESS 1
#define AHV_REASON_RING1_SIGMA 2
#define AHV_REASON_NEEDS_CONSENSUS 3
#define AHV_REASON_RING2_SIGMA 4
#define AHV_REASON_RING_INSUFFICIENT 5

#define AHV_EXPIRY_NONE INT64_MAX /* VouchRecord.expiry is None (liability/vouching.py:30) */

typedef struct ahv_ctx ahv_ctx;

/* ------------------------------------------------------------------ context */
int ahv_abi_version(void);
const char *ahv_last_error(void);
/* One context per GPU per process.  device = CUDA ordinal. */
int ahv_create(int device, ahv_ctx **out_ctx);
void ahv_destroy(ahv_ctx *ctx);
/* Waits for everything the context has enqueued.  Also reports (once) a sanity flag a kernel of a _dev entry
 * point raised since the last call, as AHV_E_INVALID: today that is a max_session_leaves smaller than a
 * session's real size (the affected roots are zeros). */
int ahv_sync(ahv_ctx *ctx);
/* sm_count, max SM clock (kHz), total device memory (bytes); any pointer may be NULL */
int ahv_device_info(ahv_ctx *ctx, int *sm_count, int *sm_clock_khz, uint64_t *total_mem);
/* cudaStream_t of the context, as void* (for callers that enqueue their own work) */
void *ahv_stream(ahv_ctx *ctx);
/* number of kernels this context has launched since creation */
uint64_t ahv_launch_count(ahv_ctx *ctx);

/* memory helpers so that callers without a CUDA binding can keep arenas resident */
int ahv_dev_alloc(ahv_ctx *ctx, uint64_t bytes, void **out_dptr);
int ahv_dev_free(ahv_ctx *ctx, void *dptr);
int ahv_host_alloc(ahv_ctx *ctx, uint64_t bytes, void **out_hptr); /* pinned */
/* flags: AHV_HOST_WRITE_COMBINED = upload-only buffer (fast for the GPU to read, very slow for the CPU to read back) */
#define AHV_HOST_WRITE_COMBINED 1u
int ahv_host_alloc_flags(ahv_ctx *ctx, uint64_t bytes, uint32_t flags, void **out_hptr);
int ahv_host_free(ahv_ctx *ctx, void *hptr);
int ahv_memcpy_h2d(ahv_ctx *ctx, void *dptr, const void *hptr, uint64_t bytes); /* async on ctx stream */
int ahv_memcpy_d2h(ahv_ctx *ctx, void *hptr, const void *dptr, uint64_t bytes); /* async on ctx stream */
int ahv_memset_dev(ahv_ctx *ctx, void *dptr, int value, uint64_t bytes);
/* Upload on the context's SECOND stream, so that it overlaps the kernels already enqueued on the main one
 * (double-buffered pipelines); ahv_upload_fence makes everything enqueued on the main stream afterwards wait for
 * the uploads issued so far.  The caller keeps source and destination untouched until then. */
int ahv_upload_async(ahv_ctx *ctx, void *dptr, const void *hptr, uint64_t bytes);
int ahv_upload_fence(ahv_ctx *ctx);

/* CUDA-event timers on the ctx stream (slot 0..63): start/stop bracket work, elapsed in ms */
int ahv_timer_start(ahv_ctx *ctx, int slot);
int ahv_timer_stop(ahv_ctx *ctx, int slot);
int ahv_timer_elapsed_ms(ahv_ctx *ctx, int slot, float *out_ms); /* synchronises on the stop event */
/* Waits until everything enqueued on the ctx stream BEFORE ahv_timer_stop(slot) has finished (work enqueued after it may
 * still be running): a completion marker for pipelines that keep the GPU fed.  Reports sticky device flags like ahv_sync. */
int ahv_timer_wait(ahv_ctx *ctx, int slot);

/* ------------------------------------------------------------------ audit path
 *
 * ahv_sha256_batch: digest[i] = SHA-256(message i).
 *   replaces SemanticDelta.compute_hash's hashlib call, audit/delta.py:63,
 *   applied to n pre-images the host serialised per audit/delta.py:43-62.
 *   offsets may be NULL when fixed_len > 0: message i = arena[i*fixed_len ..).
 */
int ahv_sha256_batch(ahv_ctx *ctx, const uint8_t *arena, const uint64_t *offsets, uint64_t n,
                     uint64_t fixed_len, uint8_t *out_digests /* n*32 */);
/*
 * ahv_sha256_one: the same for ONE message, for the call-at-a-time users of the reference's API
 * (DeltaEngine.capture -> compute_hash, audit/delta.py:85-115,63).  A message of up to AHV_ONE_MAX bytes is
 * passed to the device inside the kernel launch and its digest comes back through mapped host memory: one
 * launch and one synchronisation, no copies.  Longer messages take the ahv_sha256_batch path.
 */
#define AHV_ONE_MAX 3584u
int ahv_sha256_one(ahv_ctx *ctx, const uint8_t *msg, uint64_t len, uint8_t *out_digest /* 32 */);
int ahv_sha256_batch_dev(ahv_ctx *ctx, const uint8_t *d_arena, uint64_t arena_bytes,
                         const uint64_t *d_offsets, uint64_t n, uint64_t fixed_len,
                         uint8_t *d_out_digests);

/*
 * ahv_merkle_roots: one Merkle root per session over its leaf digests.
 *   replaces DeltaEngine.compute_merkle_root, audit/delta.py:117-134:
 *   node = SHA-256(hex(left) || hex(right)) (128 ASCII bytes), odd level
 *   duplicates its last node (:129), one leaf -> root is the leaf (:125),
 *   no leaves -> None (:119-120), reported as out_empty[s] = 1 and a zero root.
 *   max_session_leaves: the largest session size (0 = let the library find it;
 *   the _dev variant then synchronises once).  It sizes the in-CTA finish of the tree: if it is given too
 *   small, the sessions that do not fit get zero roots and the next ahv_sync returns AHV_E_INVALID.
 */
int ahv_merkle_roots(ahv_ctx *ctx, const uint8_t *digests /* n*32 */, const uint64_t *session_rowptr,
                     uint64_t n_sessions, uint8_t *out_roots /* S*32 */, uint8_t *out_empty /* S */);
int ahv_merkle_roots_dev(ahv_ctx *ctx, const uint8_t *d_digests, const uint64_t *d_session_rowptr,
                         uint64_t n_sessions, uint64_t n_leaves, uint64_t max_session_leaves,
                         uint8_t *d_out_roots, uint8_t *d_out_empty);

/*
 * ahv_merkle_roots_noalloc_dev: ahv_merkle_roots_dev for callers that must not allocate device memory
 * while running (a GPU shared with another framework's allocator under a fixed memory budget; loops
 * that cannot afford the device-wide synchronisation of cudaMalloc/cudaFree).  No counterpart in the
 * reference.  Its scratch is one fixed 8 MiB array inside the library image, held by every GPU that loads
 * the library until the process exits.  Batches needing more (about 32 bytes per leaf + 72 bytes per
 * session; roughly 250 000 leaves) are refused with AHV_E_TOOLARGE; max_session_leaves must be > 0.
 * The first call on a device looks up the array's address and should come before any stream capture on
 * that device.  Calls share the array (nothing in it needs clearing between calls), so calls that
 * overlap on the GPU race: issue them one after another.
 */
int ahv_merkle_roots_noalloc_dev(ahv_ctx *ctx, const uint8_t *d_digests, const uint64_t *d_session_rowptr,
                                 uint64_t n_sessions, uint64_t n_leaves, uint64_t max_session_leaves,
                                 uint8_t *d_out_roots, uint8_t *d_out_empty);

/*
 * ahv_commit: fused leaf hashing + Merkle roots (what Hypervisor.terminate_session
 * needs at core.py:204 for many sessions at once): audit/delta.py:63 then :117-134.
 *   out_digests may be NULL.  uniform_leaves > 0 states that every session has
 *   exactly that many leaves (session_rowptr may then be NULL).
 */
int ahv_commit(ahv_ctx *ctx, const uint8_t *arena, const uint64_t *offsets, uint64_t fixed_len,
               const uint64_t *session_rowptr, uint64_t n_sessions, uint64_t uniform_leaves,
               uint8_t *out_digests, uint8_t *out_roots, uint8_t *out_empty);
int ahv_commit_dev(ahv_ctx *ctx, const uint8_t *d_arena, uint64_t arena_bytes, const uint64_t *d_offsets,
                   uint64_t fixed_len, const uint64_t *d_session_rowptr, uint64_t n_sessions,
                   uint64_t n_leaves, uint64_t uniform_leaves, uint64_t max_session_leaves,
                   uint8_t *d_out_digests, uint8_t *d_out_roots, uint8_t *d_out_empty);

/*
 * ahv_verify_chains: DeltaEngine.verify_chain for many sessions, audit/delta.py:136-152.
 *   Effective rule of the reference (compute_hash at :140 overwrites delta_hash, so
 *   the test at :141 cannot fail): the first delta's parent must be None (:145-147)
 *   and every later parent must equal the hash recomputed for the delta before it
 *   (:149).  out_verdict[s] = 1/0, out_first_bad[s] = index inside the session of the
 *   first failing delta or -1.  out_digests (optional) receives every recomputed
 *   leaf digest (the reference writes those back into delta_hash for indices
 *   <= first_bad).  If expected_roots is given, the Merkle root is rebuilt from the
 *   recomputed digests and compared on the device (CommitmentEngine.verify,
 *   audit/commitment.py:59-64): out_root_match[s] = 1/0.  out_roots optional.
 */
int ahv_verify_chains(ahv_ctx *ctx, const uint8_t *arena, const uint64_t *offsets, uint64_t fixed_len,
                      const uint8_t *parent_digests /* n*32 */, const uint8_t *parent_kind /* n */,
                      const uint64_t *session_rowptr, uint64_t n_sessions, uint64_t uniform_leaves,
                      const uint8_t *expected_roots /* S*32 or NULL */, uint8_t *out_verdict,
                      int64_t *out_first_bad, uint8_t *out_digests, uint8_t *out_roots,
                      uint8_t *out_root_match);
int ahv_verify_chains_dev(ahv_ctx *ctx, const uint8_t *d_arena, uint64_t arena_bytes,
                          const uint64_t *d_offsets, uint64_t fixed_len, const uint8_t *d_parent_digests,
                          const uint8_t *d_parent_kind, const uint64_t *d_session_rowptr,
                          uint64_t n_sessions, uint64_t n_leaves, uint64_t uniform_leaves,
                          uint64_t max_session_leaves, const uint8_t *d_expected_roots,
                          uint8_t *d_out_verdict, int64_t *d_out_first_bad, uint8_t *d_out_digests,
                          uint8_t *d_out_roots, uint8_t *d_out_root_match);

/*
 * ahv_audit_sessions: ahv_commit followed by ahv_verify_chains, on a single upload of the batch (what an
 * auditor does when sessions arrive from the host: terminate_session's root, core.py:204, then verify_chain,
 * audit/delta.py:136-152, then CommitmentEngine.verify, audit/commitment.py:59-64).  Both device passes run
 * in full (the verify pass re-hashes).  expected_roots (S*32, or NULL): the STORED commitments to compare the
 * rebuilt roots with (an archive's CommitmentRecord.merkle_root per session); with NULL the rebuilt roots
 * are compared with the ones this same call has just committed, which checks the two passes against each
 * other and nothing else.  Batches of
 * at least 64 sessions and 8 MB are pipelined in session chunks (copy-in / hash / copy-out overlap), for
 * fixed-size and variable-length deltas, uniform and ragged sessions alike; pass pinned host buffers
 * (ahv_host_alloc) for that overlap to happen.  Host pointers only.
 * The claimed parents come either as the side arrays (parent_digests + parent_kind) or, with both NULL,
 * as parent_pos[n]: the byte position, inside each pre-image, of the parent_hash value's first character
 * (UINT32_MAX where the value is null); the library then reads the parents out of the pre-images on the
 * device (ahv_extract_parents_dev), which cuts the upload from 33 to 4 bytes per delta.
 */
int ahv_audit_sessions(ahv_ctx *ctx, const uint8_t *arena, const uint64_t *offsets, uint64_t fixed_len,
                       const uint8_t *parent_digests, const uint8_t *parent_kind, const uint32_t *parent_pos,
                       const uint64_t *session_rowptr, uint64_t n_sessions, uint64_t uniform_leaves,
                       const uint8_t *expected_roots /* S*32 or NULL */, uint8_t *out_digests, uint8_t *out_roots,
                       uint8_t *out_empty, uint8_t *out_verdict, int64_t *out_first_bad, uint8_t *out_root_match);

/*
 * ahv_extract_parents_dev: fills parent_digests / parent_kind from the pre-images themselves: the value
 * of "parent_hash" as serialised at audit/delta.py:61 (exactly 64 lowercase hex characters between quotes
 * -> AHV_PARENT_HEX; position UINT32_MAX, i.e. null -> AHV_PARENT_NONE; anything else -> AHV_PARENT_OTHER).
 */
int ahv_extract_parents_dev(ahv_ctx *ctx, const uint8_t *d_arena, uint64_t arena_bytes, const uint64_t *d_offsets,
                            uint64_t fixed_len, const uint32_t *d_parent_pos, uint64_t n,
                            uint8_t *d_out_parent_digests, uint8_t *d_out_parent_kind);

/*
 * ahv_chain_capture: session-parallel replay of DeltaEngine.capture's chaining,
 * audit/delta.py:101-114: for each session, in order, hash delta i, write its 64 hex
 * chars into delta i+1's pre-image at byte hexpos[i+1] (the "parent_hash" value the
 * host left as a 64-char placeholder), then continue.  hexpos[i] = UINT32_MAX for a
 * delta whose parent is null.  The arena is patched in place; digests and the parent
 * side arrays (for ahv_verify_chains) are produced.  While the call runs, out_digests also
 * holds intermediate hash states (option "chain_prefix"); every slot ends as its delta's digest.
 */
int ahv_chain_capture(ahv_ctx *ctx, uint8_t *arena, const uint64_t *offsets, const uint32_t *hexpos,
                      const uint64_t *session_rowptr, uint64_t n_sessions, uint8_t *out_digests,
                      uint8_t *out_parent_digests, uint8_t *out_parent_kind);
int ahv_chain_capture_dev(ahv_ctx *ctx, uint8_t *d_arena, uint64_t arena_bytes, const uint64_t *d_offsets,
                          const uint32_t *d_hexpos, const uint64_t *d_session_rowptr, uint64_t n_sessions,
                          uint64_t n_leaves, uint8_t *d_out_digests, uint8_t *d_out_parent_digests,
                          uint8_t *d_out_parent_kind);

/*
 * ahv_pack_deltas: host-side canonical packer.  Writes, for n deltas given as columns, exactly the bytes
 * SemanticDelta.compute_hash serialises (audit/delta.py:43-62: json.dumps({...}, sort_keys=True).encode(),
 * CPython's default separators and ensure_ascii escaping, timestamp.isoformat()) into a packed arena,
 * multi-threaded on the host.  Strings are UTF-8 blobs with offsets[count+1].  Call with out_arena = NULL to
 * get sizes: out_offsets[n+1] and *out_bytes are filled either way; out_parent_pos, when given, receives
 * the parent_pos array ahv_audit_sessions accepts in place of the parent side arrays.  Runs on the CPU by design (the host
 * packs, the GPU hashes); it needs no ahv_ctx.
 */
typedef struct ahv_delta_columns {
    uint64_t n;
    const uint8_t *delta_id_blob;   const uint64_t *delta_id_off;     /* n+1 */
    const uint8_t *session_id_blob; const uint64_t *session_id_off;   /* n+1 */
    const uint8_t *agent_did_blob;  const uint64_t *agent_did_off;    /* n+1 */
    const int64_t *turn_id;                                           /* n */
    const int64_t *timestamp_us;            /* n: microseconds since the epoch, UTC */
    const int32_t *timestamp_utcoffset_s;   /* n: tzinfo offset in seconds, INT32_MIN for a naive datetime */
    const uint8_t *parent_kind;             /* n: AHV_PARENT_* */
    const uint8_t *parent_digest;           /* n*32, used where kind == AHV_PARENT_HEX */
    const uint8_t *parent_other_blob; const uint64_t *parent_other_off;   /* n+1, used where kind == AHV_PARENT_OTHER */
    const uint64_t *change_rowptr;          /* n+1: changes of delta i are rows [rowptr[i], rowptr[i+1]) below */
    const uint8_t *change_path_blob;    const uint64_t *change_path_off;
    const uint8_t *change_op_blob;      const uint64_t *change_op_off;
    const uint8_t *change_content_blob; const uint64_t *change_content_off; const uint8_t *change_content_null;
    const uint8_t *change_prev_blob;    const uint64_t *change_prev_off;    const uint8_t *change_prev_null;
} ahv_delta_columns;

#define AHV_PACK_OFFSETS_VALID 1u   /* out_offsets (and out_parent_pos) already hold this batch's size query: only write */
int ahv_pack_deltas(const ahv_delta_columns *cols, uint64_t *out_offsets /* n+1 */, uint8_t *out_arena,
                    uint64_t arena_capacity, uint64_t *out_bytes, uint32_t *out_parent_pos /* n, or NULL */,
                    uint32_t flags);

/*
 * ahv_pack_deltas_dev: the same packer on the device (same emitter source, byte-identical output), one thread per
 * delta: a replay can upload the COLUMNS (the struct stays on the host, every pointer in it is a device pointer)
 * and never serialise on the host.  Asynchronous on the ctx stream: sizes -> offsets (d_out_offsets[n+1]) -> bytes
 * into d_out_arena (skip the write with d_out_arena = NULL).  If the pre-images would not fit arena_capacity
 * nothing is written and the next ahv_sync returns AHV_E_TOOLARGE; with out_bytes (a HOST pointer) the call
 * synchronises once, returns the byte count there and reports the overflow itself.
 */
int ahv_pack_deltas_dev(ahv_ctx *ctx, const ahv_delta_columns *cols, uint64_t *d_out_offsets /* n+1 */,
                        uint8_t *d_out_arena, uint64_t arena_capacity, uint64_t *out_bytes /* host, or NULL */,
                        uint32_t *d_out_parent_pos /* n, or NULL */, uint32_t flags);

/* ------------------------------------------------------------------ multi-GPU (one process per GPU)
 *
 * Sessions shard across GPUs with no data-path exchange; the one exchange step collects every
 * rank's roots and verdicts.  The communicator (NCCL, opened at run time from libnccl.so.2 or the path in
 * AHV_NCCL_LIB) belongs to the context.  No counterpart in the reference: it is a single-process library;
 * the call these stand in for is the loop that would visit every session's CommitmentEngine record
 * (audit/commitment.py:42-64) once sessions live on several devices.
 *
 *   ahv_comm_unique_id   rank 0 makes the 128-byte id; the launcher hands it to the other ranks (file, env, ...)
 *   ahv_comm_init        every rank, same id: collective, returns when all `world` ranks have joined.
 *                        world == 1 needs no id and no NCCL.
 *   ahv_comm_all_gather(_dev)  recv[r*bytes .. (r+1)*bytes) = rank r's send buffer; _dev is asynchronous on
 *                        the ctx stream, the host form stages, gathers, copies back and synchronises.
 *   ahv_gather_results_dev  ONE all-gather per step: rank r's block of `ahv_gather_stride(S)` bytes in d_all
 *                        holds its S roots (32*S bytes) followed by its S verdict bytes (zero padded to a
 *                        multiple of 16); d_all has world * stride bytes.  d_verdict may be NULL (zeros).
 */
#define AHV_COMM_ID_BYTES 128
int ahv_comm_unique_id(uint8_t *out_id /* AHV_COMM_ID_BYTES */);
int ahv_comm_init(ahv_ctx *ctx, int world, int rank, const uint8_t *id /* AHV_COMM_ID_BYTES */);
int ahv_comm_destroy(ahv_ctx *ctx);
int ahv_comm_info(ahv_ctx *ctx, int *world, int *rank, int *nccl_version); /* any pointer may be NULL */
int ahv_comm_all_gather_dev(ahv_ctx *ctx, const void *d_send, void *d_recv, uint64_t bytes_per_rank);
int ahv_comm_all_gather(ahv_ctx *ctx, const void *send, void *recv, uint64_t bytes_per_rank);
uint64_t ahv_gather_stride(uint64_t n_sessions);
int ahv_gather_results_dev(ahv_ctx *ctx, const uint8_t *d_roots /* S*32 */, const uint8_t *d_verdict /* S or NULL */,
                           uint64_t n_sessions, uint8_t *d_all /* world * ahv_gather_stride(S) */);

/* ------------------------------------------------------------------ liability / rings
 *
 * ahv_sigma_eff_rings: VouchingEngine.compute_sigma_eff for every agent of one session
 * (liability/vouching.py:128-151) over a CSR of vouch records grouped by vouchee in
 * insertion order (:186-197), then ExecutionRing.from_sigma_eff (models.py:34-42) and
 * RingEnforcer.should_demote (rings/enforcer.py:134-137).
 *   record e contributes iff active[e] && !(expiry[e] != AHV_EXPIRY_NONE && now > expiry[e])
 *   (vouching.py:34-38; times in microseconds since the epoch).
 *   sum is sequential fp64 from 0.0; sigma_eff = sigma + (omega * sum), two roundings;
 *   result = (1.0 < x) ? 1.0 : x  (Python min(), NaN kept).
 *   Rows of more than 64 records (option "sigma_long_cut") are summed in a fixed parallel shape instead
 *   (by one warp: 32 strided partials, then a fixed butterfly: deterministic, within a few ulp of the
 *   sequential sum; the run contract allows 1e-6 relative); option "sigma_exact" = 1 keeps the sequential order for every row.
 *   out_near[a] = 1 when sigma_eff lies within 1e-6 (relative) of 0.60 or 0.95: the agents whose ring could
 *   differ from the reference's under that tolerance.
 */
int ahv_sigma_eff_rings(ahv_ctx *ctx, const uint32_t *rowptr /* A+1 */, const double *bonded,
                        const uint8_t *active, const int64_t *expiry, int64_t now, const double *sigma,
                        const double *omega, const uint8_t *consensus, const uint8_t *current_ring,
                        uint64_t n_agents, double *out_sigma_eff, uint8_t *out_ring, uint8_t *out_demote,
                        uint8_t *out_near);
int ahv_sigma_eff_rings_dev(ahv_ctx *ctx, const uint32_t *d_rowptr, const double *d_bonded,
                            const uint8_t *d_active, const int64_t *d_expiry, int64_t now,
                            const double *d_sigma, const double *d_omega, const uint8_t *d_consensus,
                            const uint8_t *d_current_ring, uint64_t n_agents, double *d_out_sigma_eff,
                            uint8_t *d_out_ring, uint8_t *d_out_demote, uint8_t *d_out_near);

/*
 * ahv_exposure_sums: per-row sequential fp64 sum of the live bonded amounts of a CSR (rows = vouchers,
 * out-edges in insertion order): VouchingEngine.get_total_exposure (liability/vouching.py:157-166) and
 * LiabilityMatrix.total_exposure (liability/__init__.py:64-66) for every voucher at once.  Same edge
 * predicate as ahv_sigma_eff_rings; no cap.  Both reference methods use Python's built-in sum(), whose
 * float arithmetic depends on the interpreter: compensated = 1 reproduces CPython >= 3.12 (Neumaier
 * compensated summation, Python/bltinmodule.c), compensated = 0 the plain left-to-right additions of
 * CPython <= 3.11.  The Python shim passes the flag of the interpreter it runs under.
 */
int ahv_exposure_sums(ahv_ctx *ctx, const uint32_t *rowptr /* R+1 */, const double *bonded, const uint8_t *active,
                      const int64_t *expiry, int64_t now, uint64_t n_rows, int compensated, double *out_sum);
int ahv_exposure_sums_dev(ahv_ctx *ctx, const uint32_t *d_rowptr, const double *d_bonded, const uint8_t *d_active,
                          const int64_t *d_expiry, int64_t now, uint64_t n_rows, int compensated,
                          double *d_out_sum);

/*
 * ahv_ring_check: RingEnforcer.check for n (agent, action) pairs, rings/enforcer.py:44-128,
 * with ActionDescriptor.required_ring (models.py:122-132) derived from action_flags.
 */
int ahv_ring_check(ahv_ctx *ctx, const uint8_t *agent_ring, const uint8_t *action_flags,
                   const double *sigma_eff, const uint8_t *consensus, const uint8_t *witness, uint64_t n,
                   uint8_t *out_allowed, uint8_t *out_reason, uint8_t *out_required);
int ahv_ring_check_dev(ahv_ctx *ctx, const uint8_t *d_agent_ring, const uint8_t *d_action_flags,
                       const double *d_sigma_eff, const uint8_t *d_consensus, const uint8_t *d_witness,
                       uint64_t n, uint8_t *d_out_allowed, uint8_t *d_out_reason, uint8_t *d_out_required);

/*
 * ahv_slash_cascade: SlashingEngine.slash (liability/slashing.py:62-141) for many sessions in one launch:
 * the vouchee's score goes to 0.0 (:88); every live vouch record of its row, in insertion order
 * (get_vouchers_for, vouching.py:153-155), clips its voucher to max(score * (1.0 - risk_weight), 0.05)
 * (:94-96, SIGMA_FLOOR :54; Python max(): a NaN stays) and is released (:108, active[e] = 0); then, while
 * depth < MAX_CASCADE_DEPTH = 2 (:53,:122), each of those clips with sigma_after < 0.05 + 0.01 whose
 * voucher still has a live record is slashed one level deeper, depth first, same risk_weight (:123-138).
 *
 * Graph: the CSR the sigma_eff pass uses (rows = agents, records grouped by vouchee in insertion
 * order, live = active && !(expiry != AHV_EXPIRY_NONE && now > expiry)) plus voucher[e] = the ROW INDEX of
 * record e's voucher.  Session s owns the agent rows session_agent_rowptr[s] .. [s+1] (an agent present in
 * two sessions has two rows) and the events event_rowptr[s] .. [s+1], applied in order to the same state;
 * sessions are independent (one warp each).  scores is the dense agent_scores dict (a missing key is
 * 0.0, :94), updated in place; touched[a] = 1 where the reference would have written the key.
 *
 * Logs, in the order the reference appends to _slash_history (:119, parent before its cascades):
 * session s's slash records start at slash_log[event_rowptr[s] + rowptr[a_lo]] (slash_count[s] of them;
 * capacity n_events + E in total) and its clip records at clip_log[rowptr[a_lo]] (clip_count[s]; capacity
 * E: a record is clipped at most once because the clip releases it).  status[s]: bit 0 = an event named an
 * agent outside the session, bit 1 = a record's voucher lies outside the session (both skipped).
 * The reason strings, slash ids and timestamps of SlashResult are made on the host from these records.
 */
typedef struct ahv_slash_record {   /* one SlashingEngine.slash call (SlashResult, slashing.py:18-31) */
    uint32_t agent;                 /* row of the vouchee */
    int32_t parent;                 /* index, within the session's records, of the slash this cascaded from; -1 = an event */
    uint32_t event;                 /* index into event_agent[] of the event it belongs to */
    uint32_t depth;                 /* cascade_depth */
    uint32_t clip_begin;            /* its clips: clip records clip_begin .. +clip_count of the session */
    uint32_t clip_count;
    double sigma_before;            /* vouchee_sigma_before: the event's value, or the clip's sigma_after */
} ahv_slash_record;                 /* 32 bytes */

typedef struct ahv_clip_record {    /* VoucherClip, slashing.py:34-42 */
    uint32_t edge;                  /* index of the vouch record (gives vouch_id and voucher) */
    uint32_t slash;                 /* index, within the session's slash records, of the slash that made it */
    double sigma_before;
    double sigma_after;
} ahv_clip_record;                  /* 24 bytes */

typedef struct ahv_slash_batch {
    uint64_t n_agents;
    const uint32_t *rowptr;                 /* n_agents + 1 */
    const uint32_t *voucher;                /* E */
    uint8_t *active;                        /* E, in/out */
    const int64_t *expiry;                  /* E */
    int64_t now;
    uint64_t n_sessions;
    const uint64_t *session_agent_rowptr;   /* n_sessions + 1, last = n_agents */
    const uint64_t *event_rowptr;           /* n_sessions + 1 */
    const uint32_t *event_agent;            /* n_events: row of the violating agent */
    const double *event_sigma;              /* n_events: vouchee_sigma (recorded only) */
    const double *event_omega;              /* n_events: risk_weight */
    const uint8_t *event_depth;             /* n_events or NULL (= 0): the cascade_depth argument of slash(), :70 */
    double *scores;                         /* n_agents, in/out */
    uint8_t *touched;                       /* n_agents, out */
    ahv_slash_record *slash_log;            /* n_events + E, out */
    ahv_clip_record *clip_log;              /* E, out */
    uint32_t *slash_count;                  /* n_sessions, out */
    uint32_t *clip_count;                   /* n_sessions, out */
    uint8_t *status;                        /* n_sessions, out */
} ahv_slash_batch;

int ahv_slash_cascade(ahv_ctx *ctx, const ahv_slash_batch *batch);       /* every pointer in *batch is a host pointer */
int ahv_slash_cascade_dev(ahv_ctx *ctx, const ahv_slash_batch *batch);   /* ... a device pointer (the struct itself on the host) */

/* ------------------------------------------------------------------ measurement
 * ahv_int_peak: register-only kernel issuing the SHA-256 instruction mix
 * (funnel shifts : 3-input logic : 3-input adds = 6:4:4 per round) on every SM;
 * returns the measured rate in 32-bit integer thread-instructions per second.
 * No counterpart in the reference (it is the roofline denominator of SURVEY 8d).
 *   mode 0: ALU-pipe only (SHF/LOP3/IADD3)   mode 1: adds issued as IMAD on the FMA pipe
 */
int ahv_int_peak(ahv_ctx *ctx, int mode, int iters, double *out_instr_per_s, float *out_ms);

/* Selects a kernel variant for experiments (bench/tuning only); unknown keys return AHV_E_INVALID.
 *   "sort" 0/1          length-sort variable-length batches (default 1)
 *   "block" 64/128/256  threads per block of the leaf kernels (default 128)
 *   "fixed_variant" 0/1/2 fixed-size leaf kernel: direct 16-byte loads (default) / cp.async staged / TMA bulk-copy staged
 *   "compress" 0/1      plain additions / scaled-sum compression with additions on the FMA pipe (default 1)
 *   "chain_lanes" 0..32 chained capture: sessions per warp (0, default: as few as keep the warps within the number of
 *                   warp schedulers - thinner warps touch fewer lines per load, but two warps on one scheduler take turns)
 *   "chain_prefix" 0/1/2  chained capture: the blocks in front of each delta's parent field are hashed in a parallel
 *                   pre-pass: never / for up to 2^16 sessions, where the length of a chain bounds the pass (default 1) /
 *                   always (2).  Needs out_digests, where the intermediate states are parked
 *   "top_max" 0,2..256  level width at which a session's tree is finished in one warp/CTA (0 = auto)
 *   "blocking_sync" 0/1 ahv_audit_sessions: spin (default 0, lowest latency) / sleep on an event while its pipeline
 *                   drains, leaving the host cores to a packer thread (1; what ReplayPipeline sets)
 *   "slash_variant" 0/1/2 slashing: lane per small session + warp per larger one (default 0) / warp per session on
 *                   global memory only (1) / warp per session, staged in shared memory where it fits (2)
 *   "slash_slice" 0,1024..10752  slashing, lane kernel: bytes of shared memory per warp (multiple of 16).  0 (default):
 *                   ahv_slash_cascade picks 7040, 9216 or 10752 from the batch's average session size; ahv_slash_cascade_dev,
 *                   which does not know the record count, uses 10752 (set 7040 for sessions above ~16 agents x 48 records)
 *   "sigma_variant" 0/1/2 sigma_eff: one row per thread from global memory / tiled through shared memory (default 1) /
 *                   persistent CTAs with a two-stage pipeline (2; measured slower, kept for comparison)
 *   "sigma_exact" 0/1   sigma_eff: 1 = every row summed left to right like the reference (bit-identical, but one very long
 *                   row then bounds the pass); 0 (default) = rows longer than "sigma_long_cut" edges are summed by a
 *                   warp in a fixed shape that depends on the row alone (a few ulp from the reference, same result every run)
 *   "sigma_long_cut" 4..  that row length (default 64)
 *   "sigma_rows" 0,1..128  sigma_eff: agent rows per tile of the tiled kernel (a power of two).  0 (default): 128, or
 *                   fewer when ahv_sigma_eff_rings sees rows of more than 16 records on average (a tile then stays within
 *                   about one staging chunk); ahv_sigma_eff_rings_dev does not know the record count and uses 128
 *   "sigma_share" 0/1   CTAs of that kernel take over tiles with many more edges than a chunk from a shared queue (default 1)
 *   "sigma_heavy" 256..  a tile of that kernel with more edges than this is handed out through the shared queue (default 3072)
 *   "sigma_chunk"   records per shared-memory chunk of that kernel: 1024 / 1296 / 1440 / 1536 / 2048 (default 1440: eight CTAs per SM)
 *   "leaf_timer" 0|base record CUDA-event timers base..base+15 around every leaf-kernel launch */
int ahv_set_option(ahv_ctx *ctx, const char *key, int64_t value);

#ifdef __cplusplus
}
#endif
#endif /* AHV_H */
