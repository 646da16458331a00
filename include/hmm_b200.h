/* hmm_b200.h — C ABI of libhmm_b200.so: B200 (sm_100a) kernels for the hmm_layer forward / forward-backward path.
 *
 * Conventions (all entry points):
 *   - every pointer is a DEVICE pointer unless stated otherwise; tensors are contiguous float32, row-major;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); calls are stream-ordered and work on
 *     the CURRENT device (cudaSetDevice before the call); one process may drive several devices from several threads:
 *     the only library state is per device (raised shared-memory limits, SM count, profiling events) and mutex-guarded;
 *     the tuning / profiling switches are process-wide; inputs are never written (the reference's emitter mutates
 *     its input, defect D5);
 *   - nothing is allocated inside: the caller owns inputs, outputs and the workspace whose size the matching
 *     *_workspace_size call reports;
 *   - return value 0 = success, negative = HMM_B200_E*; hmm_b200_last_error() gives a thread-local message.
 *
 * "reference" below = sukui-genomics-cn/hmm_layer; file:line cite the interface each entry point replaces.
 */
#ifndef HMM_B200_H
#define HMM_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HMM_B200_OK 0
#define HMM_B200_EINVAL (-1)     /* bad argument (shape, null pointer, alignment) */
#define HMM_B200_EWORKSPACE (-2) /* workspace too small */
#define HMM_B200_ECUDA (-3)      /* CUDA runtime error, see hmm_b200_last_error() */
#define HMM_B200_EUNSUPPORTED (-4)

/* flags */
#define HMM_B200_TRAINING 1u  /* emitter `training=True`: +1e-7 on the codon factor (gene_pred_hmm_emitter.py:260-261) */
#define HMM_B200_NO_LOGLIK 2u /* posterior without the "- loglik" (MsaHMMLayer.py:513-514) */
#define HMM_B200_ALPHA_SCRATCH 4u /* dense_run: out_alpha is only lent as scratch, its final contents are not needed */
#define HMM_B200_BETA_SCRATCH 8u  /* dense_run: same for out_beta */
#define HMM_B200_ALPHA_NORMALIZED 16u /* genepred_run: out_alpha = log of the per-position normalised forward vector
                                        (the scaled state of MsaHmmCell.py:92) instead of log alpha; input of genepred_grad */

int hmm_b200_abi_version(void);
const char* hmm_b200_last_error(void);

/* Optional tuning of the chunked scan: chunk length in positions (0 = chosen per call from the problem size and
 * the SM count; at most 40), threads per CTA of the finalize kernel (0 = derived; a multiple of 64 in [128,512]).
 * dense_path selects the step kernel of the generic dense path: 0 = by size (tcgen05 tensor-core step from 192 states),
 * 1 = CUDA-core step always, 2 = tensor-core step always.  Process-wide; meant for benchmarking and tests. */
int hmm_b200_set_tuning(int chunk_len, int finalize_threads, int dense_path);

/* Posterior-only calls of hmm_b200_genepred_run take a single-launch kernel whenever every tile of the call can be
 * resident on the device at once (gene15_fused.cu); mode 1 switches that off (two-kernel path: products + finalize),
 * mode 0 (default) switches it back on.  Process-wide; meant for benchmarking and tests. */
int hmm_b200_set_fused_mode(int mode);

/* Per-kernel timing for benchmarks: when enabled, hmm_b200_genepred_run records CUDA events around each of its
 * launches (products, [boundary], finalize); hmm_b200_profile_read waits for the last profiled call and writes up to n
 * durations in milliseconds to the HOST array ms_host, returning how many it wrote. */
int hmm_b200_profile_enable(int on);
/* Debug: when device_buffer is non-NULL, thread 0 of every CTA of the gene-HMM kernels writes clock64() phase
 * timestamps into it (int64 [2][CTAs][16]); NULL switches it off (tools/phase_times.py). */
int hmm_b200_debug_phases(void* device_buffer);
int hmm_b200_profile_read(float* ms_host, int n);

/* ---------------------------------------------------------------------------------------------------------
 * 15-state gene-structure HMM (Tiberius), emission fused.  Replaces, for GenePredHMMEmitter +
 * GenePred(Multi)HMMTransitioner(k=1) cells:
 *   _state_posterior_log_probs_impl  MsaHMMLayer.py:422-521   (out_post)
 *   _forward_recursion_impl          MsaHMMLayer.py:227-282   (out_alpha, loglik)
 *   _backward_recursion_impl         MsaHMMLayer.py:322-381   (out_beta)
 *   HmmCell.emission_probs           MsaHmmCell.py:61-71  -> GenePredHMMEmitter.forward gene_pred_hmm_emitter.py:231-277
 *   HmmCell.forward / BaseRNN loop   MsaHmmCell.py:73-106, BaseRNN.py:216-227
 *   Bidirectional.forward            Bidirectional.py:113-164
 *
 *   inputs    (num_models, batch, T, 20)  15 class probabilities ‖ one-hot A,C,G,T,N
 *   A         (num_models, 15, 15)        transition matrix; only the 23 gene-model edges are read
 *                                         (gene_pred_hmm_transitioner.py:200-221)
 *   pi        (num_models, 15)            initial distribution
 *   Bfull     (num_models, 15, 15)        [state][class] = softmax(emission kernel), intron rows already repeated
 *   codon5    (2, 125, 12)                5-symbol (A,C,G,T,N) left/right 3-mer tables for states E2,START,EI0-2,
 *                                         IE0-2,STOP (+3 pad), built from the emitter's (2,9,64) codon_probs by
 *                                         hmm_b200_genepred_prepare_codon (once per parameter set)
 *   end_hints (num_models, batch, 2, 15)  or NULL
 *   out_post / out_alpha / out_beta (num_models, batch, T, 15), each may be NULL
 *   loglik    (num_models, batch)         may be NULL
 * --------------------------------------------------------------------------------------------------------- */
size_t hmm_b200_genepred_workspace_size(int num_models, int batch, int T);

/* codon64 (2,9,64) = GenePredHMMEmitter.codon_probs (gene_pred_hmm_emitter.py:198-217)  ->  codon5 (2,125,12):
 * index (near*5 + far)*5 + pivot over A,C,G,T,N where N (and positions outside the sequence) average the four
 * bases, which is what the reference's soft k-mers do for N (kmer.py:21-33). */
int hmm_b200_genepred_prepare_codon(const float* codon64, float* codon5, void* stream);

int hmm_b200_genepred_run(const float* inputs, int num_models, int batch, int T,
                          const float* A, const float* pi, const float* Bfull, const float* codon5,
                          const float* end_hints, unsigned flags,
                          float* out_post, float* out_alpha, float* out_beta, float* loglik,
                          void* workspace, size_t workspace_bytes, void* stream);

/* The same evaluation with HOST buffers (what a caller holding CPU tensors binds: the reference takes CPU tensors and
 * returns CPU tensors, MsaHMMLayer.py:134-153).  host_inputs / host_end_hints / host_out_* / host_loglik are HOST
 * pointers (pinned memory for full copy/compute overlap; pageable memory works, serialised by the driver); A, pi, Bfull,
 * codon5 stay DEVICE pointers (parameters live on the GPU).  The batch is cut into slices: slice i's kernels run on
 * `stream` while slice i+1's inputs and slice i-1's outputs move over PCIe on two internal copy streams (created once per
 * thread and device; no device memory is allocated inside: device_scratch is the caller's, sized by
 * hmm_b200_genepred_host_scratch_size).  The call returns when everything is queued; the outputs are complete once the
 * work queued on `stream` up to this call has finished (cudaStreamSynchronize(stream)). */
size_t hmm_b200_genepred_host_scratch_size(int num_models, int batch, int T);

int hmm_b200_genepred_run_host(const float* host_inputs, int num_models, int batch, int T,
                               const float* A, const float* pi, const float* Bfull, const float* codon5,
                               const float* host_end_hints, unsigned flags,
                               float* host_out_post, float* host_out_alpha, float* host_out_beta, float* host_loglik,
                               void* device_scratch, size_t device_scratch_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Compact wire format of the same evaluation.  The reference's layout — 15 fp32 class probabilities
 * followed by a 5-wide fp32 one-hot of the nucleotide (gene_pred_hmm_emitter.py:239-240, 80 B / position) — is dictated
 * by its emitter's slicing; callers that produce the class probabilities in bf16 (the upstream network's dtype) and
 * hold the sequence as bytes hand over 31 B / position instead:
 *   class_probs_bf16 (num_models, batch, T, 15)  bf16 bit patterns
 *   nuc_codes        (num_models, batch, T)      0..3 = A,C,G,T, anything else = N (1/4 on each base, kmer.py:21-25)
 * Results are those of hmm_b200_genepred_run on the fp32 expansion of exactly these values.  Workspace / scratch sizes
 * are the ones of the fp32 entry points.  Sequence sharding is not offered in this format.
 * --------------------------------------------------------------------------------------------------------- */
int hmm_b200_genepred_run_compact(const unsigned short* class_probs_bf16, const unsigned char* nuc_codes,
                                  int num_models, int batch, int T,
                                  const float* A, const float* pi, const float* Bfull, const float* codon5,
                                  const float* end_hints, unsigned flags,
                                  float* out_post, float* out_alpha, float* out_beta, float* loglik,
                                  void* workspace, size_t workspace_bytes, void* stream);

int hmm_b200_genepred_run_host_compact(const unsigned short* host_class_probs_bf16, const unsigned char* host_nuc_codes,
                                       int num_models, int batch, int T,
                                       const float* A, const float* pi, const float* Bfull, const float* codon5,
                                       const float* host_end_hints, unsigned flags,
                                       float* host_out_post, float* host_out_alpha, float* host_out_beta, float* host_loglik,
                                       void* device_scratch, size_t device_scratch_bytes, void* stream);

/* Stand-alone emission (cell.emission_probs API parity): out_E (num_models, batch, T, 15). */
int hmm_b200_genepred_emission(const float* inputs, int num_models, int batch, int T,
                               const float* Bfull, const float* codon5, const float* end_hints,
                               unsigned flags, float* out_E, void* stream);

/* Stand-alone emission of the gene-structure emitters for any number of classes / states / gene-model copies:
 *   has_codon = 0: reference SimpleGenePredHMMEmitter.forward (gene_pred_hmm_emitter.py:93-121), inputs
 *                  (num_models, batch, T, num_classes), E = x . Bfull^T, end hints on the first / last position;
 *   has_codon = 1: reference GenePredHMMEmitter.forward (:231-263) with num_copies >= 1, inputs
 *                  (num_models, batch, T, num_classes + 5), num_states == 1 + 14 * num_copies, codon columns repeated per
 *                  copy (:254-257), 1/4096 for the 1 + 5 * num_copies unconstrained states, +1e-7 with HMM_B200_TRAINING.
 * Bfull (num_models, num_states, num_classes): one softmax row per state (intron rows repeated, :115-116).  codon5: the
 * (2,125,12) tables of hmm_b200_genepred_prepare_codon (NULL when has_codon = 0).  out_E (num_models, batch, T, num_states).
 * Model i reads inputs[i] (the shipped reference reads inputs[0] for every model, defect D13). */
int hmm_b200_emission_general(const float* inputs, int num_models, int batch, int T, int num_classes, int num_states,
                              int num_copies, int has_codon, const float* Bfull, const float* codon5,
                              const float* end_hints, unsigned flags, float* out_E, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Gradient of the gene-HMM log-likelihood (training through the layer).  The reference has no backward code of its
 * own: autograd differentiates its Python time loop (HmmCell.forward MsaHmmCell.py:73-106, the emitter
 * gene_pred_hmm_emitter.py:93-121,231-263, parameters rebuilt per call :65,80 and gene_pred_hmm_transitioner.py:66-71).
 * This entry point replaces that traced backward pass with the closed form (expected transition / emission counts):
 *   log_post       (num_models, batch, T, 15)  out_post of hmm_b200_genepred_run (without HMM_B200_NO_LOGLIK)
 *   log_alpha_hat  (num_models, batch, T, 15)  out_alpha of the same call made with HMM_B200_ALPHA_NORMALIZED
 *   grad_loglik    (num_models, batch)         upstream gradient of every log-likelihood
 *   grad_inputs    (num_models, batch, T, 20)  or NULL; the 5 nucleotide flags are data and get zeros
 *   grad_A         (num_models, 15, 15)        or NULL; non-zero on the 23 gene-model edges only
 *   grad_pi        (num_models, 15)            or NULL
 *   grad_Bfull     (num_models, 15, 15)        or NULL; [state][class]
 * flags: HMM_B200_TRAINING as in the forward call.  Deterministic (fixed-order partial sums, no atomics).
 * --------------------------------------------------------------------------------------------------------- */
size_t hmm_b200_genepred_grad_workspace_size(int num_models, int batch, int T);

int hmm_b200_genepred_grad(const float* inputs, int num_models, int batch, int T,
                           const float* A, const float* pi, const float* Bfull, const float* codon5,
                           const float* end_hints, unsigned flags,
                           const float* log_post, const float* log_alpha_hat, const float* grad_loglik,
                           float* grad_inputs, float* grad_A, float* grad_pi, float* grad_Bfull,
                           void* workspace, size_t workspace_bytes, void* stream);

/* Parameter construction of the 15-state gene model in one launch, and its backward (a training step rebuilds A, pi, B
 * from the kernels on every call: GenePredHMMTransitioner.make_A / make_initial_distribution
 * gene_pred_hmm_transitioner.py:66-112 through make_transition_matrix_from_indices Transitioner.py:337-380, and
 * SimpleGenePredHMMEmitter.make_B gene_pred_hmm_emitter.py:87-91 with the intron row repeated :115-116).
 *   transition_kernel (23)  in the edge order of GenePredHMMTransitioner.make_transition_indices (k = 1)
 *   start_kernel (15)       emission_kernel (num_models, 13, 15)
 *   -> A (num_models,15,15), pi (num_models,15), Bfull (num_models,15,15).  grad_* inputs of the backward may be NULL. */
int hmm_b200_genepred_params(const float* transition_kernel, const float* start_kernel, const float* emission_kernel,
                             int num_models, float* A, float* pi, float* Bfull, void* stream);
int hmm_b200_genepred_params_grad(const float* transition_kernel, const float* start_kernel, const float* emission_kernel,
                                  int num_models, const float* grad_A, const float* grad_pi, const float* grad_Bfull,
                                  float* grad_transition_kernel, float* grad_start_kernel, float* grad_emission_kernel,
                                  void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Viterbi decoding of the gene HMM: most probable state path.  The reference names Viterbi (MsaHmmCell.py:13) and
 * prepares log_A_dense for it (:46-47) but ships no implementation; this is the max-product form of the recursion
 * MsaHmmCell.py:73-106 over the same floored emissions / initial distribution, ties to the lowest state index.
 *   path   (num_models, batch, T) uint8 state indices      score (num_models, batch) log joint of the path, or NULL
 * The log emissions are materialised in the workspace.  Sequences of 1024 to 131072 positions are decoded by a chunked
 * (max,+) scan (64 chunks per sequence: chunk transfer matrices, a fold, the recursion inside every chunk, back-pointers in
 * shared memory; a cluster of 4 CTAs per sequence), shorter or longer ones by one warp per sequence (T dependent steps).
 * A must carry the 23-edge topology of the single-copy gene model (entries outside it are ignored).
 * --------------------------------------------------------------------------------------------------------- */
size_t hmm_b200_genepred_viterbi_workspace_size(int num_models, int batch, int T);

int hmm_b200_genepred_viterbi(const float* inputs, int num_models, int batch, int T,
                              const float* A, const float* pi, const float* Bfull, const float* codon5,
                              const float* end_hints, unsigned flags,
                              unsigned char* path, float* score,
                              void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Sequence sharding over GPUs (one genome-length sequence cut into contiguous shards, one per rank).  Replaces
 * the reference's chunk combine with the chunks living on different devices:
 *   _get_total_forward_from_chunks / _get_total_backward_from_chunks   MsaHMMLayer.py:285-319, 384-419
 *   TotalProbabilityCell.forward                                       TotalProbabilityCell.py:30-49
 *
 * This rank holds positions [seq_off, seq_off + T_local) of every sequence of total length seq_total.
 *   inputs     (num_models, batch, T_local, 20)
 *   halo       (num_models*batch, 4, 5)  nucleotide one-hot rows of positions seq_off-2, seq_off-1, seq_off+T_local,
 *                                        seq_off+T_local+1 (rows outside the sequence are ignored); the 3-mer windows
 *                                        reach two positions across a shard edge (gene_pred_hmm_emitter.py:245-249)
 *   end_hints  apply to positions 0 and seq_total-1 of the whole sequence
 * Step 1, every rank:  shard_reduce  -> block_out (num_models*batch, 256): the shard's transfer matrix as one
 *                      stored block (scaled rows, row scales, power-of-two exponent).
 * Step 2, caller:      ONE all-gather of block_out over the ranks -> all_blocks (world, num_models*batch, 256).
 * Step 3, every rank:  shard_finalize with the SAME workspace (it holds the shard's product trees from step 1)
 *                      -> outputs for the local positions; loglik (num_models, batch) of the whole sequence.
 * --------------------------------------------------------------------------------------------------------- */
int hmm_b200_genepred_shard_reduce(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                   const float* halo, const float* A, const float* pi, const float* Bfull,
                                   const float* codon5, const float* end_hints, unsigned flags, float* block_out,
                                   void* workspace, size_t workspace_bytes, void* stream);

int hmm_b200_genepred_shard_finalize(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                     const float* halo, const float* A, const float* pi, const float* Bfull,
                                     const float* codon5, const float* end_hints, unsigned flags, const float* all_blocks,
                                     int rank, int world, float* out_post, float* out_alpha, float* out_beta, float* loglik,
                                     void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Generic dense-transition HMM (any q: profile HMMs q = 2L+3, multi-copy gene models q = 1+14k) from a
 * materialised emission tensor.  Replaces, for cells whose plugins are not the fused 15-state gene model:
 *   HmmCell.forward + BaseRNN loop     MsaHmmCell.py:73-106, BaseRNN.py:216-227
 *   Transitioner.__call__ (x.A, x.A^T) Transitioner.py:320-331, gene_pred_hmm_transitioner.py:114-125
 *   _forward_recursion_impl / _backward_recursion_impl / _state_posterior_log_probs_impl (parallel_factor = 1
 *   semantics)                         MsaHMMLayer.py:227-282, 322-381, 422-521
 *
 *   E            (num_models, batch, T, q)   emission probabilities (cell.emission_probs)
 *   A            (num_models, q, q)          transition matrix;  A_transposed: its transpose (backward pass), may be
 *                                            NULL when neither out_beta nor out_post is requested
 *   pi           (num_models, q)
 *   out_alpha / out_beta (num_models, batch, T, q): log alpha / log beta; out_post needs BOTH buffers (they hold
 *                the per-step vectors; pass scratch for the one you do not want back)
 *   loglik       (num_models, batch), may be NULL
 *   flags        HMM_B200_NO_LOGLIK, HMM_B200_ALPHA_SCRATCH, HMM_B200_BETA_SCRATCH (the scratch flags skip
 *                writing log alpha / log beta back into lent buffers: 40 % less HBM traffic in the finishing pass)
 * --------------------------------------------------------------------------------------------------------- */
size_t hmm_b200_dense_workspace_size(int num_models, int batch, int T, int q);

int hmm_b200_dense_run(const float* E, const float* A, const float* A_transposed, const float* pi,
                       int num_models, int batch, int T, int q, unsigned flags,
                       float* out_post, float* out_alpha, float* out_beta, float* loglik,
                       void* workspace, size_t workspace_bytes, void* stream);

/* The same evaluation with the emission contraction fused for one-hot inputs (reference Emitter.py:139-157: inputs . B^T
 * picks a column of B when the input row is one-hot): E is never materialised.
 *   symbols (num_models, batch, T) uint8     the symbol of every position (values >= num_symbols are clamped)
 *   Bt      (num_models, num_symbols, q)     Bt[m][x][j] = P(symbol x | state j) = B[m][j][x]
 * The persistent recurrence (q < 192) stages a sequence's symbols in shared memory and fills its emission ring from rows
 * of the (L2-resident) table; the tcgen05 step (q >= 192) gathers the 32 rows of an epilogue warp by symbol.  Shapes
 * no gathering engine covers expand the tensor inside the workspace (hmm_b200_dense_workspace_size_onehot accounts for it). */
size_t hmm_b200_dense_workspace_size_onehot(int num_models, int batch, int T, int q);

int hmm_b200_dense_run_onehot(const unsigned char* symbols, const float* Bt, int num_symbols,
                              const float* A, const float* A_transposed, const float* pi,
                              int num_models, int batch, int T, int q, unsigned flags,
                              float* out_post, float* out_alpha, float* out_beta, float* loglik,
                              void* workspace, size_t workspace_bytes, void* stream);

/* One HmmCell.forward call (MsaHmmCell.py:73-106), for API parity with code that drives the cell step by step.
 *   E_t (num_models*n, q); scaled (num_models*n, qc*q); loglik (num_models*n, qc), qc = 1 or q (chunked form);
 *   A (num_models, q, q) already oriented (A forward, A^T reverse); init = 1 skips the transition (:77-78)
 *   out (num_models*n, qc*q + qc) = [log a-hat | loglik] forward, [log R | old loglik] reverse (:96-103) */
int hmm_b200_cell_step(const float* E_t, const float* scaled, const float* loglik, const float* A,
                       int num_models, int n, int qc, int q, int reverse, int init,
                       float* out, float* new_scaled, float* new_loglik, void* stream);

/* The same three steps with the exchange INSIDE the kernels (peer memory over NVLink) and the boundary machinery folded
 * into two kernels: 4 launches per step (products, shard product + peer stores, wait + cross-rank / cross-group fold +
 * tile replay, finalize) instead of 7 launches around an NCCL all-gather.
 *   peer_block_tables / peer_flag_tables: HOST arrays of `world` device pointers, the block table (float [2][world][NB][256])
 *     and flag table (uint32 [2][world][NB]) of every rank in rank order, all mapped into this process (CUDA IPC /
 *     torch symmetric memory); the tables must be zero before the first step.  NULL: no peer stores -- the block goes to
 *     block_out and the caller gathers it (e.g. NCCL), then passes the gathered blocks with block_flags = NULL.
 *   step_tag: step counter, identical on every rank, never 0; its parity selects the table half, so a rank may run one
 *     step ahead of its peers (it cannot run two ahead: finalize_x waits for every rank's block of the step).
 *   finalize_x: all_blocks / block_flags = this rank's OWN tables at half (step_tag & 1).
 * The workspace must be zero before the first step (a per-sequence counter lives there and is left at zero). */
int hmm_b200_genepred_shard_reduce_x(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                     const float* halo, const float* A, const float* pi, const float* Bfull, const float* codon5,
                                     const float* end_hints, unsigned flags, float* block_out,
                                     const unsigned long long* peer_block_tables, const unsigned long long* peer_flag_tables,
                                     int rank, int world, unsigned step_tag, void* workspace, size_t workspace_bytes, void* stream);

int hmm_b200_genepred_shard_finalize_x(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                       const float* halo, const float* A, const float* pi, const float* Bfull, const float* codon5,
                                       const float* end_hints, unsigned flags, const float* all_blocks, const unsigned* block_flags,
                                       int rank, int world, unsigned step_tag, float* out_post, float* out_alpha, float* out_beta,
                                       float* loglik, void* workspace, size_t workspace_bytes, void* stream);

/* Sequence sharding in ONE launch per step (the single-launch kernel of hmm_b200_genepred_run with reducer CTAs): every
 * tile of the shard computes emission, chunk matrices and its tile matrix; one extra CTA per 32 tiles multiplies them,
 * one more multiplies those and stores the shard's block into every rank's table (peer stores over NVLink, tagged flags as
 * shard_reduce_x), all of them wait for the other ranks' blocks and fold ranks -> groups -> tiles (the role of
 * TotalProbabilityCell.py:30-49 across devices) while the tile CTAs wait inside the kernel; then recursion and output
 * as in the unsharded kernel.  Posterior (+ log-likelihood) only, peer tables required; interchangeable rank by rank
 * with shard_reduce_x + shard_finalize_x of the same step (same tables, same tags).
 * hmm_b200_genepred_shard_fused_chunk: the chunk length the launch would use, 0 when the shard does not fit one launch
 * on the current device (all of its CTAs must be resident at once) -- then shard_run_x fails with HMM_B200_EINVAL. */
int hmm_b200_genepred_shard_fused_chunk(int num_models, int batch, int T_local, int world);

int hmm_b200_genepred_shard_run_x(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                  const float* halo, const float* A, const float* pi, const float* Bfull, const float* codon5,
                                  const float* end_hints, unsigned flags, const unsigned long long* peer_block_tables,
                                  const unsigned long long* peer_flag_tables, int rank, int world, unsigned step_tag,
                                  float* out_post, float* loglik, void* workspace, size_t workspace_bytes, void* stream);

/* hmm_b200_genepred_run without a workspace argument and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, outputs and results as hmm_b200_genepred_run; the scratch
 * is a static device array inside the library (HMM_B200_NOALLOC_SCRATCH_BYTES, held by every GPU that loads the library
 * until the process exits), whose address is resolved once per device.  The call itself issues no cudaMalloc / cudaFree
 * or allocating library routine.
 * Limit: hmm_b200_genepred_workspace_size(num_models, batch, T) <= HMM_B200_NOALLOC_SCRATCH_BYTES (about 110 000
 * positions in total); larger problems fail with HMM_B200_EWORKSPACE and a message in hmm_b200_last_error().
 * Every call on a device shares the one array: calls that overlap on the GPU (two streams, two threads) race and give
 * wrong results -- serialise them (one stream, or events).  The array keeps the previous call's contents; the call
 * clears what has to start at zero on `stream`. */
#define HMM_B200_NOALLOC_SCRATCH_BYTES (64u << 20)
int hmm_b200_genepred_run_noalloc(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi,
                                  const float* Bfull, const float* codon5, const float* end_hints, unsigned flags,
                                  float* out_post, float* out_alpha, float* out_beta, float* loglik, void* stream);

#ifdef __cplusplus
}
#endif
#endif
