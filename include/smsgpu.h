/* smsgpu.h -- C ABI of libsmsgpu.so, the B200-native odd-cycle separation engine for SMS.
 *
 * The reference (CharJon/SMS, paths relative to its root) has NO C ABI for this path: the SCIP constraint
 * handler `cycleparity` owns a C++ object sms::OcwSeparator by value (include/sms/conshdlr/conshdlr_oscw.hpp:24-25)
 * and calls its methods directly.  This header is the thin C layer BASELINE.json's north_star asks for;
 * every entry point names the reference interface it replaces.  The reference-side binding a maintainer
 * adds (a GpuOcwSeparator with the public surface of sms::OcwSeparator) is shown in INTEGRATION.md and
 * shipped in sms_b200/host/.
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every function returns 0 on success and a
 * negative smsgpu_status otherwise (no exception crosses the ABI); smsgpu_last_error() gives the text.
 * A context is single-threaded (SCIP calls its plugins from one thread: src/solver/basic_ocw.cpp:239,242).
 * There is NO CPU fallback: without a CUDA device smsgpu_create fails with SMSGPU_ERR_CUDA.
 *
 * Numbering contract: vertices 0..n-1 compact; adjacency of a vertex in
 * NetworKit::Graph::addEdge insertion order (extern/networkit/networkit/cpp/graph/Graph.cpp:472-520);
 * edge ids as assigned by Graph::indexEdges (Graph.cpp:131-170).  "doubled id" = 2*vertex + side, the
 * double-cover node numbering of include/sms/graph/ocw_seperator.hpp:66-79.
 */
#ifndef SMSGPU_H
#define SMSGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SMSGPU_ABI_VERSION 3

typedef struct smsgpu_ctx smsgpu_ctx;

typedef enum smsgpu_status {
    SMSGPU_OK = 0,
    SMSGPU_ERR_ARG = -1,      /* bad argument (null pointer, non-compact graph, self loop, multi-edge ...);
                                 the reference throws std::runtime_error here (src/graph/ocw_seperator.cpp:16-19) */
    SMSGPU_ERR_CUDA = -2,     /* CUDA runtime / driver error or no device */
    SMSGPU_ERR_NOMEM = -3,    /* device or host allocation failed */
    SMSGPU_ERR_STATE = -4,    /* call order violated (e.g. separate before set_round_input) */
    SMSGPU_ERR_INTERNAL = -5, /* a source could not be finished in any tier / dedup self-check failed */
    SMSGPU_ERR_NCCL = -6      /* NCCL not loadable or a collective failed */
} smsgpu_status;

/* Parameters of one separation round.  smsgpu_default_params fills the reference's hard-coded values. */
typedef struct smsgpu_params {
    int32_t max_intermediates; /* 10: "first 10 intermediates", src/graph/ocw_seperator.cpp:343 */
    double best_dist_init;     /* 1.0 - 1e-6: initial bestDist, ocw_seperator.cpp:51 */
    int32_t dedup;             /* 1: canonical rotation/direction + sort + unique (K5).  0: raw walk list in
                                  (source index, emission order), exactly what getViolatedOcws returns per source */
    int32_t export_labels;     /* 1: keep every stored Dijkstra label (node, dist, hops) per source -- parity tests only */
    int32_t export_src_stats;  /* 1: per-source (arcs scanned, settled) counters */
    int32_t start_tier;        /* 0 = auto (shared-memory tier first), 1 / 2 = HBM hashed tiers (medium / large), 3 = dense tier */
    int32_t mode;              /* SMSGPU_MODE_INDEPENDENT (default) or SMSGPU_MODE_VERBATIM, see below */
    int32_t debug_flags;       /* 0; debugging aids of the search kernels (tests only) */
    double min_efficacy;       /* 0.0 = keep everything.  Cuts with violation / sqrt(len) <= min_efficacy are dropped on the
                                  device before dedup: the SCIPisCutEfficacious test of conshdlr_oscw.cpp:221,232 (efficacy =
                                  (activity - rhs) / ||a||_2 = (1 - length) / sqrt(len); separating/minefficacyroot = 1e-6,
                                  src/solver/basic_ocw.cpp:244) as a PRE-filter with a relative slack of 1e-9; the handler
                                  still asks SCIP for the survivors */
    int32_t reserved[6];
} smsgpu_params;

/* The reference's sweep over the sources is ORDER DEPENDENT: a source that finds nothing blacklists its vertex and
 * everything reachable over integral edges (blacklistIntReachable, ocw_seperator.cpp:113-114,363-382), and building a
 * walk blacklists the stem it cuts off (:162-169); later sources skip blacklisted neighbours (:98-99) and blacklisted
 * sources (:334-335).
 *   SMSGPU_MODE_INDEPENDENT  every source sees the round's INITIAL blacklist (what sharding sources over GPUs means);
 *                            the cut set is a superset-like variant of the reference's (DESIGN.md section 2 has the
 *                            measured differences).
 *   SMSGPU_MODE_VERBATIM     the cut list of the reference's own sequential sweep, reproduced exactly: the blacklist is
 *                            kept as a per-vertex TIME STAMP (index of the first source that blacklists the vertex), every
 *                            source reads "blacklisted before me", and the round is iterated to the fixed point of the
 *                            stamps (smsgpu_stats.verbatim_sweeps sweeps).  Single GPU.  Cost: several
 *                            sweeps per round -- far ahead of the reference's CPU sweep on fractional x*, but on tie-heavy LP
 *                            iterates (long dependency chains between consecutive sources, ~100 sweeps) BEHIND it; measured
 *                            numbers in DESIGN.md section 5. */
#define SMSGPU_MODE_INDEPENDENT 0
#define SMSGPU_MODE_VERBATIM 1

/* Host-visible result of one round; all pointers are owned by the context and stay valid until the next
 * smsgpu_separate / smsgpu_destroy.  Cut k is the closed walk
 *   node[off[k]] -e0- node[off[k]+1] -e1- ... -e(len-1)- node[off[k]]
 * with eid[off[k]+i] / is_stay[off[k]+i] describing the edge LEAVING node[off[k]+i]
 * (OddClosedWalk::getIthNode / ithEdgeIsStay, include/sms/graph/odd_closed_walk.hpp:43-51).
 * The SCIP row is  sum_{cross} x_e - sum_{stay} x_e <= n_cross - 1  (src/conshdlr/conshdlr_oscw.cpp:199-216). */
typedef struct smsgpu_cuts {
    int64_t ncuts;
    int64_t nnodes;          /* total length of node / eid / is_stay */
    const int64_t *off;      /* ncuts + 1 */
    const int32_t *node;     /* original vertex ids */
    const int32_t *eid;      /* NetworKit edge ids -> ProbDataRootedMc::edgeIdToVar (mc_rooted.hpp:71) */
    const uint8_t *is_stay;  /* 1 = coefficient -1, 0 = cross edge, coefficient +1 */
    const int32_t *n_cross;  /* per cut: numCrosses(), rhs = n_cross - 1 */
    const int32_t *src;      /* per cut: index into the source list (dedup: the smallest index that found it) */
    const double *violation; /* per cut: 1 - length  (> 0 means violated by x*); efficacy = violation / sqrt(len) */
} smsgpu_cuts;

typedef struct smsgpu_stats {
    int64_t sources_run;       /* sources not skipped by the blacklist test (ocw_seperator.cpp:334) */
    int64_t arcs_scanned;      /* == numberOfEdgesSeen summed over sources (ocw_seperator.cpp:86,100) */
    int64_t settled;           /* heap extractions summed over sources */
    int64_t raw_walks;         /* walks before dedup */
    int64_t unique_cuts;
    int64_t tier_sources[4];   /* sources finished in the shared-memory / HBM medium / HBM large / dense tier */
    double ms_weights;         /* K1 */
    double ms_search;          /* K2-K4 fused: bounded shortest paths + walk extraction + chord repair */
    double ms_dedup;           /* K5 */
    double ms_total_device;    /* first kernel start -> last kernel end */
    double ms_h2d, ms_d2h;
    int64_t search_launches;   /* kernel launches of the search kernel this round */
    int64_t total_launches;    /* all kernel launches this round */
    int64_t mirrored_sources;  /* sources whose results were derived from their twin's search (same vertex, other side) */
    int64_t mirrored_arcs;     /* part of arcs_scanned that was NOT scanned: the twin's scans stand for them */
    double ms_tier[4];         /* search kernel time per tier (shared / HBM medium / HBM large / dense); the HBM large
                                * tier may run beside the medium one, so the four need not add up to ms_search */
    int64_t verbatim_sweeps;   /* SMSGPU_MODE_VERBATIM: sweeps over the sources until the stamps reached their fixed point
                                * (arcs_scanned / settled are summed over them); 0 in independent-source mode */
} smsgpu_stats;

/* ---- life cycle: replaces OcwSeparator::OcwSeparator (src/graph/ocw_seperator.cpp:9-23) ---------------------- */
/* Uploads the CSR once.  rowptr[n+1], col[2m], eid[2m] in the numbering contract above.  device = CUDA ordinal. */
int smsgpu_create(smsgpu_ctx **out, int64_t n, int64_t m, const int64_t *rowptr, const int32_t *col,
                  const int32_t *eid, int device);
void smsgpu_destroy(smsgpu_ctx *ctx);
const char *smsgpu_last_error(const smsgpu_ctx *ctx); /* ctx may be NULL: error of the last failed smsgpu_create */
int smsgpu_abi_version(void);
void smsgpu_default_params(smsgpu_params *p);

/* ---- per round input: replaces the m x updateWeights (ocw_seperator.cpp:25-35, called from
 * conshdlr_oscw.cpp:175-180) and the n x updateBlacklist (ocw_seperator.hpp:25-28, conshdlr_oscw.cpp:182-185).
 * x_by_eid[m] is the LP point indexed by edge id (clamped to [0,1] on the device); blacklist[n] is 1 for the
 * root and locally fixed vertices.  Host buffers; pageable ones are copied through pinned staging, page-locked ones
 * (the context's own, below, or the caller's cudaHostAlloc / cudaHostRegister memory) are read by the DMA where they
 * lie and must stay untouched until the round has been launched. */
int smsgpu_set_round_input(smsgpu_ctx *ctx, const double *x_by_eid, const uint8_t *blacklist);
/* The context's own pinned staging buffers (x[m], blacklist[n]).  A handler that gathers x* edge by edge
 * (SCIPgetSolVal per edge, conshdlr_oscw.cpp:175-180) writes straight into them and passes the same pointers to
 * smsgpu_set_round_input, which then skips the pageable -> pinned copy.  Valid until smsgpu_destroy. */
int smsgpu_input_buffers(smsgpu_ctx *ctx, double **x_by_eid, uint8_t **blacklist);
/* Same, with buffers already resident on ctx's device (bench "value" leg; a GPU LP solver would use this). */
int smsgpu_set_round_input_device(smsgpu_ctx *ctx, const double *d_x_by_eid, const uint8_t *d_blacklist);

/* Source list: doubled ids, in the order ConshdlrOscw issues them (conshdlr_oscw.cpp:188-195 passes
 * 0..n-1).  sources == NULL selects 0..n-1.  With (rank, nranks) the list is dealt cyclically in pairs (2p, 2p+1) and this
 * context runs only its share (source sharding across GPUs).  A context remembers, per source,
 * how much work the previous round was and issues the next round accordingly (results never depend on it); setting
 * the SAME list again keeps that memory, a different list drops it. */
int smsgpu_set_sources(smsgpu_ctx *ctx, const int32_t *sources, int64_t nsrc, int rank, int nranks);

/* ---- one separation round: replaces the loop over OcwSeparator::getViolatedOcws(source)
 * (ocw_seperator.cpp:330-361, driven by conshdlr_oscw.cpp:188-195) in independent-source mode: every source sees
 * the round's initial blacklist (DESIGN.md "mode I").  Blocking; results via smsgpu_get_cuts. */
int smsgpu_separate(smsgpu_ctx *ctx, const smsgpu_params *params);
/* Same as smsgpu_separate but leaves the result on the device (no D2H); pair with smsgpu_fetch_cuts. */
int smsgpu_separate_async(smsgpu_ctx *ctx, const smsgpu_params *params);
int smsgpu_fetch_cuts(smsgpu_ctx *ctx);
int smsgpu_get_cuts(smsgpu_ctx *ctx, smsgpu_cuts *out);
int smsgpu_get_stats(smsgpu_ctx *ctx, smsgpu_stats *out);

/* Parity instrumentation (params.export_labels / export_src_stats).  label_off[nsrc_local+1] indexes the
 * label arrays per local source index; nodes are doubled ids.  The exported set is every label the search
 * kept: all settled nodes, all queue members and all intermediates (far labels that provably cannot matter
 * are never created, DESIGN.md "label pruning"); each equals the reference's distances_/hopDistances_. */
int smsgpu_get_labels(smsgpu_ctx *ctx, int64_t *nlabels, const int64_t **label_off, const int32_t **label_node,
                      const double **label_dist, const int32_t **label_hops);
int smsgpu_get_src_stats(smsgpu_ctx *ctx, int64_t *nsrc, const int64_t **arcs, const int64_t **settled,
                         const int32_t **tier);

/* ---- small-cycle scan: replaces TriangleSeparator::getViolatedTriangles (include/sms/conshdlr/triangles.hpp:34-53)
 * and HoleSeparator::getViolatedFourHoles (include/sms/conshdlr/holes.hpp:38-54), both built on
 * violatedOddSelectionClosedWalk (src/graph/odd_closed_walk.cpp:8-65).  k = 3 or 4; cycles[k*T] are vertex ids
 * in the order addTriangle / addFourHole received them.  The scan uses the x* of the last set_round_input and
 * returns walks in list order through smsgpu_get_small_cycle_cuts (same layout as smsgpu_cuts; src = list index). */
int smsgpu_small_cycles_set(smsgpu_ctx *ctx, int k, int64_t T, const int32_t *cycles);
/* The lists themselves, computed on the device: replaces SmallChordlessCycles::computeSmallCCs (src/graph/small_ccs.cpp:18-43,
 * called from ConshdlrTriangles::scip_initsol, src/conshdlr/conshdlr_triangles.cpp:193-203, and ConshdlrHoles).  Triangles
 * {x, u, y} and (want_holes) chordless four-cycles {x, u, y, z}, in the reference's list order; both lists are registered
 * as smsgpu_small_cycles_set would.  smsgpu_small_cycles_get returns a host copy of a registered list. */
int smsgpu_small_cycles_enumerate(smsgpu_ctx *ctx, int want_holes, int64_t *ntriangles, int64_t *nholes);
int smsgpu_small_cycles_get(smsgpu_ctx *ctx, int k, int64_t *T, const int32_t **cycles);
int smsgpu_small_cycles_scan(smsgpu_ctx *ctx, int k);
int smsgpu_small_cycles_scan_device(smsgpu_ctx *ctx, int k); /* same, the walks stay on the device (bench "value" leg) */
int smsgpu_get_small_cycle_cuts(smsgpu_ctx *ctx, int k, smsgpu_cuts *out);
/* device times of the last scan (CUDA events on the context's stream): the scan kernel, the ranking + gather of the
 * violated cycles, the device-to-host copy of the walks */
int smsgpu_small_cycles_timing(smsgpu_ctx *ctx, double *ms_scan, double *ms_gather, double *ms_d2h);

/* ---- the small-cycle scan without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same result as smsgpu_small_cycles_scan (the violated cycles of the list, in list order, same arrays as
 * smsgpu_cuts with src = list index), but stateless and context free: every input and output is a DEVICE buffer of the
 * caller, and the scan's scratch (one code byte per cycle + a count and a base per 1024 cycles) is a file-scope
 * __device__ array of SMSGPU_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  The call issues
 * no cudaMalloc / cudaFree / allocating library routine: three kernel launches on `stream` (a cudaStream_t, NULL = the
 * default stream) of `device`, plus an 8-byte D2H copy and a stream synchronize only when h_count is given.
 *   k               3 or 4
 *   T               cycles in the list, 0 <= T <= SMSGPU_NOALLOC_MAX_CYCLES (larger: SMSGPU_ERR_ARG, nothing is launched)
 *   d_cycles        [k*T] vertex ids, cycle by cycle (what smsgpu_small_cycles_get returns)
 *   d_cycle_eids    [k*T] edge id of the edge (cycle[i], cycle[(i+1) % k]); every id must index d_x (not checked)
 *   d_x             x* by edge id
 *   out_capacity    violated cycles the output arrays can hold; d_node, d_eid, d_is_stay [k*out_capacity], d_n_cross, d_src,
 *                   d_violation [out_capacity], d_off [out_capacity + 1].  Violated cycles beyond it are dropped (the first
 *                   out_capacity in list order are kept), *d_count still reports all of them
 *   d_count         (device) number of violated cycles;  h_count (host, may be NULL): the same, copied back
 *   err, err_len    optional buffer for the message of a failed call
 * Every call on a device shares the scratch array, which keeps the last call's contents (nothing in it needs clearing:
 * each call writes every word it later reads).  Calls that overlap on the GPU -- two streams, two threads, two
 * processes' worth of contexts are fine, two calls in flight on one device are not -- race on it.  The first call on
 * a device resolves the array's device address (cudaGetSymbolAddress); make that call outside a graph capture. */
#define SMSGPU_NOALLOC_MAX_CYCLES 4194304
#define SMSGPU_NOALLOC_SCRATCH_BYTES 4243968
int smsgpu_small_cycles_scan_noalloc(int device, void *stream, int k, int64_t T, const int32_t *d_cycles,
                                     const int32_t *d_cycle_eids, const double *d_x, int64_t out_capacity, int32_t *d_node,
                                     int32_t *d_eid, uint8_t *d_is_stay, int32_t *d_n_cross, int32_t *d_src,
                                     double *d_violation, int64_t *d_off, int64_t *d_count, int64_t *h_count, char *err,
                                     size_t err_len);

/* ---- multi-GPU: one context per GPU (one process per GPU under torchrun, or one thread per GPU inside a
 * single SCIP process).  After smsgpu_comm_init every smsgpu_separate ends with an NCCL all-gather of the
 * per-GPU cut lists and a global dedup, so each rank returns the identical full list. */
int smsgpu_nccl_unique_id(void *id128);                                   /* rank 0; 128 bytes */
int smsgpu_comm_init(smsgpu_ctx *ctx, const void *id128, int rank, int nranks);
int smsgpu_comm_destroy(smsgpu_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* SMSGPU_H */
