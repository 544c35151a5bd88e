/*
 * herdb.h -- C ABI of the B200 gate-evaluation backend for herd-fhe/herdsman.
 *
 * This is the drop-in boundary for herdsman's one data-parallel hot path:
 * "apply a boolean circuit, gate by gate, to every row of an encrypted
 * data-frame partition".  Every entry point cites the reference interface it
 * stands behind (paths relative to the herdsman tree):
 *
 *   herdb_map_async        == IWorkerGroup::schedule_task(MapTask)
 *                             include/execution/worker/i_worker_group.hpp:44-48,
 *                             common/include/herd/common/model/worker/task.hpp:12-21,
 *                             wire form proto/src/worker.proto:28-37,57-61
 *   herdb_done_fn/status   == IWorkerGroup::TaskHandle::Status + completion callback
 *                             include/execution/worker/i_worker_group.hpp:15-42,
 *                             src/execution/worker/i_worker_group.cpp:4-20
 *   herdb_circuit/node/edge== herd::common::Circuit / node_t / DAG edges
 *                             common/include/herd/common/model/circuit/circuit.hpp:14-30,
 *                             nodes.hpp:12-43, operation.hpp:12-20, dag.hpp (parents() order
 *                             = edge insertion order, dag.hpp:208-230)
 *   herdb_load_key         == <key_dir>/<session>/<int(SchemaType)>.key
 *                             src/service/key_service.cpp:19-44
 *   partition files        == <storage_dir>/<session>/<data_frame>/<partition>, rows framed
 *                             [u32 len][payload]  src/service/storage_service.cpp:19-28,99-150
 *   gate arithmetic        == OpenFHE binfhe v1.0.x EvalBinGate/EvalNOT (not in the reference
 *                             tree; restated in oracle/oracle.c)
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every function
 * returns 0 or a negative herdb_status code and never throws; herdb_strerror()
 * names a code, herdb_last_error() gives the detailed message of the last
 * failure on that context.  There is NO CPU fallback: without a CUDA device
 * herdb_create fails with HERDB_ERR_CUDA.
 */
#ifndef HERDB_H
#define HERDB_H

#include <stddef.h>
#include <stdint.h>

#include "herdb_params.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct herdb_ctx herdb_ctx; /* opaque: devices, streams, device copies of keys */

enum herdb_status {
    HERDB_OK = 0,
    HERDB_ERR_ARG = -1,      /* bad argument / malformed circuit                   */
    HERDB_ERR_CUDA = -2,     /* CUDA runtime/driver error, or no device            */
    HERDB_ERR_IO = -3,       /* key / partition file missing or malformed          */
    HERDB_ERR_FORMAT = -4,   /* key blob / row payload does not match parameters   */
    HERDB_ERR_NOKEY = -5,    /* no key loaded for the session                      */
    HERDB_ERR_NOMEM = -6,
    HERDB_ERR_UNSUPPORTED = -7
};

/* IWorkerGroup::TaskHandle::Status, same order (i_worker_group.hpp:18-24) */
enum herdb_task_status { HERDB_TASK_COMPLETED = 0, HERDB_TASK_PENDING = 1,
                         HERDB_TASK_TIME_OUT = 2, HERDB_TASK_ERROR = 3 };

/* herd::common::DataType, same order (common/include/herd/common/model/data_type.hpp:9-22) */
enum herdb_data_type { HERDB_BIT = 0, HERDB_UINT8, HERDB_UINT16, HERDB_UINT32, HERDB_UINT64,
                       HERDB_INT8, HERDB_INT16, HERDB_INT32, HERDB_INT64 };

/* herd::common::Operation, same order (operation.hpp:12-20) */
enum herdb_operation { HERDB_OP_AND = 0, HERDB_OP_OR, HERDB_OP_NOT, HERDB_OP_XOR,
                       HERDB_OP_NAND, HERDB_OP_MUX };

enum herdb_node_kind { HERDB_NODE_INPUT = 0, HERDB_NODE_OUTPUT = 1,
                       HERDB_NODE_CONSTANT = 2, HERDB_NODE_OPERATION = 3 };

/* HERDB_NODE_FAST on an XOR node: evaluate it as ONE bootstrap with OpenFHE's BINGATE XOR_FAST
 * (2*(ct1 - ct2), gate constant 5q/8) instead of the OR(AND(a,!b), AND(!a,b)) form that v1.0.x
 * EvalBinGate(XOR) expands to.  Never set by the reference (its wire format cannot even carry XOR,
 * node.proto:25-30); an opt-in of this backend's own circuit builders. */
enum herdb_node_flags { HERDB_NODE_FAST = 1 };

/* node_t (nodes.hpp:12-43): which fields are meaningful depends on kind */
typedef struct herdb_node {
    uint8_t  kind;        /* herdb_node_kind                                        */
    uint8_t  op;          /* herdb_operation            (OPERATION)                 */
    uint8_t  const_val;   /* 0/1                        (CONSTANT)                  */
    uint8_t  flags;       /* herdb_node_flags           (OPERATION)                 */
    uint32_t tuple;       /* InputNode::tuple_index     (INPUT)                     */
    uint32_t field;       /* field_index                (INPUT, OUTPUT)             */
    uint32_t bit;         /* bit_index, LSB = 0         (INPUT, OUTPUT)             */
} herdb_node;

/* DAG edge parent -> child.  Operand order of a node = order in which its
 * incoming edges appear in this array (DAG::parents(), dag.hpp:208-230).
 * MUX operands are (select, value_if_1, value_if_0). */
typedef struct herdb_edge { uint32_t from, to; } herdb_edge;

/* Circuit (circuit.hpp:14-30): inputs[tuple][field] -> DataType, output[] -> DataType */
typedef struct herdb_circuit {
    const herdb_node* nodes;        size_t n_nodes;
    const herdb_edge* edges;        size_t n_edges;
    const uint8_t*  in_field_types;     /* DataType of every input field, tuples concatenated */
    const uint32_t* in_tuple_offsets;   /* n_tuples + 1 offsets into in_field_types            */
    size_t n_tuples;
    const uint8_t*  out_field_types;    /* DataType per output column                          */
    size_t n_out_fields;
} herdb_circuit;

typedef void (*herdb_done_fn)(void* user, int task_status /* herdb_task_status */);

/* ------------------------------------------------------------------ lifecycle */
int  herdb_create(herdb_ctx** out, const int* device_ids, int n_devices,
                  const char* storage_dir, const char* key_dir);
void herdb_destroy(herdb_ctx* ctx);
const char* herdb_strerror(int code);
const char* herdb_last_error(const herdb_ctx* ctx);
int  herdb_device_count(const herdb_ctx* ctx);
/* host threads that run tasks (default 2, HERDB_TASK_THREADS): what an IWorkerGroup built on this context
 * should answer to concurrent_workers() (i_worker_group.hpp:48; executor.cpp:96-113 keeps that many in flight) */
int  herdb_task_threads(const herdb_ctx* ctx);
/* tasks an IWorkerGroup built on this context should keep outstanding = its answer to concurrent_workers() (default 8,
 * HERDB_TASK_WINDOW; never below the task threads).  Deeper than the task threads on purpose: small map tasks that wait
 * in the queue behind one another -- same session, same circuit -- are coalesced into ONE evaluation of up to
 * HERDB_COALESCE_ROWS rows (default 4096, 0 = off), because a partition of a few hundred rows fills only a fraction of a
 * grid round per circuit level.  Every task keeps its own partition files, completion and status; if a coalesced
 * evaluation fails, its tasks are re-run one by one so that only the task at fault fails. */
int  herdb_task_window(const herdb_ctx* ctx);

/* ------------------------------------------------------------------ keys
 * load_key reads <key_dir>/<session>/<schema_type>.key (key_service.cpp:28-31);
 * load_key_mem takes the same blob from host memory; load_key_dev takes it from
 * device memory of ctx device `device` (e.g. after an NCCL broadcast done by the
 * caller, one process per GPU).  With n_devices > 1 the blob is uploaded to the
 * first device and reaches the others in one ncclBroadcast over NVLink / NVSwitch
 * (SURVEY.md 2.1 K5); every device then converts its copy at the same time.
 *
 * Tasks (herdb_map_async / herdb_reduce_async) load the key of their session themselves
 * and re-check the key FILE before every task: KeyService::add_key overwrites the file
 * when a session uploads a key again and remove_key deletes it (key_service.cpp:19-60).
 * A changed file (size / mtime / inode) is reloaded, a missing one fails the task.
 * File-backed keys no task is using are evicted least-recently-used first once the key
 * material per device exceeds HERDB_KEY_BUDGET_MB (default 32768; 324 MiB per STD128
 * session) and come back from their file on the session's next task. */
size_t herdb_key_blob_size(uint32_t n);     /* bytes of a key blob for LWE dimension n (herdb_params.h) */
int herdb_loaded_keys(herdb_ctx* ctx);      /* sessions whose key is resident right now */
int herdb_load_key(herdb_ctx* ctx, const char* session_uuid, int schema_type);
int herdb_load_key_mem(herdb_ctx* ctx, const char* session_uuid, const void* blob, size_t len);
int herdb_load_key_dev(herdb_ctx* ctx, const char* session_uuid, int device,
                       const void* d_blob, size_t len);
int herdb_unload_key(herdb_ctx* ctx, const char* session_uuid);
/* LWE dimension n of the session's key (ciphertext = n+1 u16 words) */
int herdb_key_lwe_n(herdb_ctx* ctx, const char* session_uuid, uint32_t* n_out);

/* ------------------------------------------------------------------ the task seam
 * == IWorkerGroup::schedule_task(MapTask).  Returns immediately; `done` is
 * invoked exactly once from a backend-owned host thread after the output
 * partition file has been written and renamed into place (never from inside
 * this call: executor.cpp:86-93 installs its callback after schedule_task
 * returns). */
int herdb_map_async(herdb_ctx* ctx, const char* session_uuid,
                    const char* in_df_uuid, uint32_t in_partition, uint64_t row_count,
                    const char* out_df_uuid, uint32_t out_partition,
                    int schema_type /* int(CryptoKeyPtr.schema_type): the key file is <schema_type>.key */,
                    const herdb_circuit* circuit, herdb_done_fn done, void* user);
/* == IWorkerGroup::schedule_task(ReduceTask) (task.hpp:23-32, worker.proto:39-48; built at
 * execution_service.cpp:474-513).  The rows of all input partitions, in the order given, are folded to ONE
 * output row with the binary circuit: inputs[0] and inputs[1] are two rows of the same column layout as
 * the output (InputNode::tuple_index selects which, nodes.hpp:12-21).  The reference worker is not in the
 * tree, so the folding order is this backend's own [INFERRED]: a balanced pairwise tree, round by round
 * over the current list (0,1),(2,3),...; an odd last row is carried unchanged to the end of the next
 * list.  Every round is one batched evaluation of the circuit over all pairs (the same level-batched
 * kernels as a map task).  The output partition receives exactly one row.
 * row_count of an input frame is a HINT: the reference labels the hidden frames of Policy::PARALLEL with
 * row_count = partitions although each holds the one row a first-level reduce wrote
 * (execution_service.cpp:613-618), so every input partition is read to its end and only the per-row length
 * is validated. */
typedef struct herdb_input_frame {
    const char* df_uuid;          /* InputDataFramePtr.pointer.uuid      (data_frame_ptr.hpp:15-19) */
    uint32_t    partition;        /* InputDataFramePtr.pointer.partition                            */
    uint64_t    row_count;        /* InputDataFramePtr.row_count (not trusted, see above)           */
} herdb_input_frame;
int herdb_reduce_async(herdb_ctx* ctx, const char* session_uuid,
                       const herdb_input_frame* inputs, size_t n_inputs,
                       const char* out_df_uuid, uint32_t out_partition, int schema_type,
                       const herdb_circuit* circuit, herdb_done_fn done, void* user);
/* the same fold over host rows (n_rows >= 1 rows of out-row layout in, one row out) */
int herdb_reduce_rows(herdb_ctx* ctx, const char* session_uuid, const herdb_circuit* circuit,
                      const uint16_t* rows_in, size_t n_rows, uint16_t* row_out);
/* blocks until every task scheduled so far has completed */
int herdb_wait_idle(herdb_ctx* ctx);
/* Partition-file framing used by the task path (host only, no device needed):
 * rows concatenated, each [u32 length, host endian][payload]  (storage_service.cpp:19-28,133-135).
 * read: exactly `row_count` rows of exactly `row_bytes` payload each, else HERDB_ERR_FORMAT / _IO.
 * write: to "<path>.tmp", then rename() so a partial file is never observed
 * (lambda_http_worker_group.cpp:244-259 treats existence of the file as success). */
int herdb_partition_read(const char* path, uint64_t row_count, size_t row_bytes, void* rows_out);
int herdb_partition_write(const char* path, uint64_t row_count, size_t row_bytes, const void* rows);
/* number of rows in a partition file whose rows all have `row_bytes` payload (reads to the end of the file) */
int herdb_partition_count(const char* path, size_t row_bytes, uint64_t* rows_out);
/* rows of partition k when `rows` rows are dealt to `partitions` files: rows/P + (k < rows%P)
 * (storage_service.cpp:121-129,328-331) */
uint64_t herdb_partition_rows(uint64_t rows, uint32_t partitions, uint32_t k);

/* ------------------------------------------------------------------ file-less entry points
 * Host buffers (the end-to-end path: H2D + evaluate + D2H).
 * A ciphertext is (n+1) little-endian u16: a[0..n) then b, residues mod q=1024.
 * A row is the concatenation of its bit-ciphertexts, fields in order, LSB first. */
int herdb_eval_gate_batch(herdb_ctx* ctx, const char* session_uuid, int device, int gate /* herdb_gate */,
                          const uint16_t* ct_a, const uint16_t* ct_b, uint16_t* ct_out, size_t batch);
int herdb_eval_circuit_rows(herdb_ctx* ctx, const char* session_uuid, const herdb_circuit* circuit,
                            const uint16_t* rows_in, uint16_t* rows_out, size_t n_rows);
/* A chain of map stages evaluated back to back (execution_service.cpp:524-549 creates one intermediate
 * frame per stage): the output row of stage k is the input row of stage k+1 and never leaves HBM; only the
 * first stage's rows go up and the last stage's rows come down. */
int herdb_eval_pipeline_rows(herdb_ctx* ctx, const char* session_uuid,
                             const herdb_circuit* const* circuits, size_t n_stages,
                             const uint16_t* rows_in, uint16_t* rows_out, size_t n_rows);
/* Device-resident variants (inputs/outputs already in HBM of ctx device `device`;
 * dense layout, ciphertext stride n+1 u16).  Work is enqueued on the backend's
 * stream for that device and the call returns after synchronising it. */
int herdb_eval_gate_batch_dev(herdb_ctx* ctx, const char* session_uuid, int device, int gate,
                              const uint16_t* d_ct_a, const uint16_t* d_ct_b, uint16_t* d_ct_out,
                              size_t batch);
int herdb_eval_circuit_rows_dev(herdb_ctx* ctx, const char* session_uuid, int device,
                                const herdb_circuit* circuit,
                                const uint16_t* d_rows_in, uint16_t* d_rows_out, size_t n_rows);
/* herdb_eval_gate_batch_dev with the CTA shape forced: cts_per_cta = 4 (throughput shape), 2 or 1 (the small-job
 * shapes a launch picks for tails and for batches that would leave SMs empty), 0 = the launch policy.  Measurement
 * and parity tests of the shapes; every shape gives the same ciphertexts. */
int herdb_eval_gate_batch_dev_shape(herdb_ctx* ctx, const char* session_uuid, int device, int gate,
                                    const uint16_t* d_ct_a, const uint16_t* d_ct_b, uint16_t* d_ct_out,
                                    size_t batch, int cts_per_cta);

/* herdb_eval_gate_batch_dev for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and ciphertexts as herdb_eval_gate_batch_dev; the call path contains no cudaMalloc, cudaFree,
 * event creation or other allocating routine.  The scratch the two kernels share (the extraction buffer between blind
 * rotation and key switch) is a fixed array in the library's device image: 4.66 MiB that every GPU which loads the
 * library holds until exit, enough for batch <= HERDB_NOALLOC_MAX_BATCH (2,368 gates, herdb_params.h).  A larger batch is
 * refused with HERDB_ERR_ARG and a message in herdb_last_error.  All calls on a device share that array (it keeps the
 * last call's contents, which is harmless: every word read is written first): calls of one context are serialised by
 * the library, but calls that overlap on the GPU -- two contexts on one device -- race.  Kernel times of this path are
 * not added to herdb_stats (timing events would have to be created); launches and gates are counted. */
int herdb_eval_gate_batch_dev_noalloc(herdb_ctx* ctx, const char* session_uuid, int device, int gate,
                                      const uint16_t* d_ct_a, const uint16_t* d_ct_b, uint16_t* d_ct_out,
                                      size_t batch);

/* circuit facts (host-only, no device needed): bootstrapped gates per row, levels, slots */
typedef struct herdb_circuit_info {
    uint32_t in_bits, out_bits;     /* ciphertexts per input / output row                      */
    uint32_t boot_gates;            /* bootstraps per row after lowering XOR/MUX               */
    uint32_t levels;                /* dependent bootstrap levels                              */
    uint32_t max_level_width;       /* widest level (gates)                                    */
    uint32_t slots;                 /* ciphertext slots per row after liveness reuse           */
} herdb_circuit_info;
int herdb_circuit_analyze(const herdb_circuit* circuit, herdb_circuit_info* info);

/* The level schedule a launch of `rows` rows gets on a grid that retires `wave_jobs` jobs (gates x rows) per
 * round of its persistent CTAs (SM count x 4; the eval entry points pass their own).  The reference evaluates
 * a row node by node (executor.cpp:98-146); here gates are grouped into dependency levels, and gates with
 * slack are placed on the level whose last, part-empty round they fill.  rows == 0: as-soon-as-possible
 * levels.  level_widths (may be NULL) receives min(levels, cap) gate counts. */
typedef struct herdb_schedule_info {
    uint32_t levels, max_level_width, slots;
    uint32_t policy;                /* 0 = as soon as possible kept, else the list schedule that won  */
    uint64_t rounds_asap, rounds;   /* grid rounds per launch before / after (0 without a hint)       */
} herdb_schedule_info;
int herdb_circuit_schedule(const herdb_circuit* circuit, uint64_t rows, uint32_t wave_jobs,
                           herdb_schedule_info* info, uint32_t* level_widths, size_t cap);
/* the compiled levels run on plaintext bits (in_bits / out_bits: one byte per bit of a row): a host-side check
 * of the compiler and its slot reuse.  HERDB_ERR_ARG if two gates of a level conflict on a slot. */
int herdb_circuit_eval_plain(const herdb_circuit* circuit, uint64_t rows, uint32_t wave_jobs,
                             const uint8_t* in_bits, uint8_t* out_bits);

/* ------------------------------------------------------------------ measurement hooks */
typedef struct herdb_stats {
    uint64_t kernel_launches;       /* our kernels launched since create / last reset          */
    uint64_t gates_bootstrapped;
    double   blind_rotate_ms;       /* CUDA-event time of blind-rotate launches (sum)          */
    double   keyswitch_ms;          /* CUDA-event time of key-switch launches (sum)            */
    uint64_t blind_rotate_launches, keyswitch_launches;
    uint64_t h2d_bytes, d2h_bytes;
    double   key_load_ms;           /* last key load: host blob -> kernel-ready keys on every device (wall clock) */
    double   key_bcast_ms;          /* of which the NCCL broadcast to the other devices (0 on one device)         */
} herdb_stats;
int herdb_get_stats(herdb_ctx* ctx, int device, herdb_stats* out);
/* the cudaStream_t (as void*) every kernel and copy of ctx device `device` is enqueued on, so a
 * caller can bracket work with its own CUDA events on the launching stream */
int herdb_get_stream(herdb_ctx* ctx, int device, void** stream_out);
int herdb_reset_stats(herdb_ctx* ctx);
/* raw integer-pipe peak microbenchmark (IMAD / IADD3 / IMAD.HI / IMAD.WIDE), Tops/s */
int herdb_measure_int_peak(herdb_ctx* ctx, int device, double* imad_tops, double* iadd3_tops,
                           double* imadhi_tops, double* imadwide_tops, double* mixed_tops);
/* L2 read bandwidth (GB/s): all CTAs streaming one 32 MiB buffer with 128-bit loads; the roof of keyswitch_kernel,
 * whose key rows are served by L2 */
int herdb_measure_l2_bandwidth(herdb_ctx* ctx, int device, double* gbs);

#ifdef __cplusplus
}
#endif
#endif /* HERDB_H */
