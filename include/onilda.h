/*
 * onilda.h - C ABI of libonilda.so: the B200 (sm_100a) implementation of ONI's
 * machine-learning hot path - the oni-lda-c variational-EM LDA and the post-LDA
 * suspicious-connects scoring.
 *
 * The reference has NO foreign-function interface for this path: oni-ml shells
 * out to the `lda` program and reads its text files back
 * (oni-ml/src/main/scala/org/opennetworkinsight/OniLDACWrapper.scala:83-104).
 * The drop-in boundary is therefore (1) the `lda est|inf` program built on this
 * library (open-network-insight_b200/host/lda_main.c) and (2) the entry points
 * below, which are what a cgo/JNI/ctypes binding of that path would bind.  Each
 * declaration cites the reference function it replaces (paths relative to the
 * reference checkout).  INTEGRATION.md shows the bindings.
 *
 * Conventions
 *   - every function returns 0 on success, a negative ONILDA_E* code otherwise;
 *     onilda_last_error() gives the message for the calling thread.
 *   - plain pointers and sizes only.  Unless a name ends in _device, pointers
 *     are HOST pointers owned by the caller; the library owns all device memory.
 *   - log_beta is row-major [K][V] (topic-major, the reference's
 *     lda_model.log_prob_w, lda.h:40-47); gamma is [D][K]; theta is [D][K];
 *     phi is word-major [V][K] (the transposed table the Scala scorer builds,
 *     OniLDACWrapper.scala:178-195).
 *   - a handle is bound to one CUDA device; calls on one handle are not
 *     thread-safe, distinct handles are independent.
 *   - there is no CPU fallback: every compute entry point fails with
 *     ONILDA_ECUDA when no CUDA device is usable.
 */
#ifndef ONILDA_H
#define ONILDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ONILDA_OK 0
#define ONILDA_EINVAL (-1)   /* bad argument */
#define ONILDA_EIO (-2)      /* file could not be opened / written */
#define ONILDA_ECUDA (-3)    /* CUDA runtime error or no device */
#define ONILDA_ENOMEM (-4)   /* host or device allocation failed */
#define ONILDA_ENAN (-5)     /* NaN likelihood (the reference asserts: lda-inference.c:78) */

typedef struct onilda_corpus onilda_corpus;
typedef struct onilda_model onilda_model;
typedef struct onilda_em onilda_em;
typedef struct onilda_scorer onilda_scorer;

/* The five values of the settings file.  Types as in the reference: both
 * tolerances are `float` (lda-inference.h:10, lda-estimate.h:21). */
typedef struct {
    int var_max_iter;     /* -1 = until converged (lda-inference.c:46-47) */
    float var_converged;
    int em_max_iter;
    float em_converged;
    int estimate_alpha;   /* 1 unless the file says "alpha fixed" */
} onilda_settings;

const char* onilda_last_error(void);
const char* onilda_version(void);
/* Number of usable CUDA devices (0 when none; never fails). */
int onilda_device_count(void);
/* Creates the CUDA contexts of the first ndev devices (<= 0: all) ahead of time.  No reference counterpart: it
 * exists so that a host program can overlap context creation with reading its input (host/lda_main.c does). */
int onilda_device_warmup(int ndev);

/* ------------------------------------------------------------------ settings
 * replaces read_settings (oni-ml/oni-lda-c/lda-estimate.c:331-350): five literal
 * lines "var max iter %d" / "var convergence %f" / "em max iter %d" /
 * "em convergence %f" / "alpha %s". */
int onilda_settings_read(const char* path, onilda_settings* out);

/* -------------------------------------------------------------------- corpus
 * replaces read_data / document / corpus (lda-data.c:22-59, lda.h:23-37).
 * Token-based, newline-agnostic "%10d" then "%10d:%10d" per entry; duplicates
 * kept in file order; num_terms = max word id + 1; per-document total is the
 * int sum of counts.  The result is a CSR corpus on the host. */
int onilda_corpus_from_dat(const char* path, onilda_corpus** out);
/* Same corpus from caller CSR arrays (copied).  num_terms <= 0 derives it from
 * the data as read_data does; a larger value is kept (vocabulary with unseen ids). */
int onilda_corpus_from_csr(const int64_t* offsets, const int32_t* words, const int32_t* counts,
                           int64_t num_docs, int32_t num_terms, onilda_corpus** out);
/* SURVEY section 8(f) row N1.  Replaces the grouping half of OniLDACWrapper.createModel
 * (OniLDACWrapper.scala:146-168) together with the model.dat round trip through read_data
 * (lda-data.c:22-59): n (document id, word id, count) triples, in any order, become the CSR
 * corpus.  Documents come out in id order (the order of `distinctDocument`), the entries of a
 * document keep their input order (createModelTest: ("192.168.1.1",w0,8), ("10.10.98.123",w1,4),
 * ("66.23.45.11",w2,2), ("192.168.1.1",w3,5) -> "2 0:8 3:5", "1 1:4", "1 2:2"), repeated words
 * stay separate entries as read_data keeps them.  num_docs 0 = max id + 1; ids that never occur
 * are empty documents.  The grouping is a stable radix sort on `device` (-1: current); the host
 * arrays may be pageable.  n < 2^31. */
int onilda_corpus_from_triples(const int32_t* doc, const int32_t* word, const int32_t* count, int64_t n,
                               int64_t num_docs, int32_t num_terms, int device, onilda_corpus** out);
/* ------------------------------------------------------- flow word creation
 * netflow here, DNS below.
 *
 * onilda_quantiles_i64: Quantiles.computeQuantiles (Quantiles.scala:53-75) of an integer column:
 * cuts_out[i] = the least x in the data with (count of values <= x) / n >= thresholds[i], compared
 * in doubles as the reference does; +infinity for an empty column or a threshold above 1.  At
 * most 16 thresholds.  Values are taken as the doubles the reference converts them to (exact
 * below 2^53). */
int onilda_quantiles_i64(const int64_t* values, int64_t n, const double* thresholds, int nq, int device,
                         double* cuts_out);
int onilda_quantiles_i64_device(const int64_t* values, int64_t n, const double* thresholds, int nq,
                                double* cuts_out /* host */, void* stream);
/* FlowWordCreation.flowWordCreation (FlowWordCreation.scala:14-84): deciles of the time of day
 * hour + minute/60 + second/3600, deciles of ibyt, quintiles of ipkt (Quantiles.scala:86,94),
 * Quantiles.bin (:97-99) of the three, and adjustPort (:92-146) per event.  The source and
 * destination words come back as integer codes
 *   code = (((port_index * 10 + timeBin) * 10 + ibytBin) * 5 + ipktBin) * 2 + marker
 * port_index = the word port (0..65535), 65536 for 111111, 65537 for 333333; marker 1 = the
 * word carries the "-1_" prefix.  onilda_flow_word_string prints a code as the reference's word
 * string ("-1_23.0_7.0_7.0_4.0").  The ip pair string of adjustPort is not part of any word and
 * is not produced.  hour 0..23, minute/second 0..59, ports 0..65535, else ONILDA_EINVAL. */
typedef struct onilda_flow_cuts {
    double time[10];
    double ibyt[10];
    double ipkt[5];
} onilda_flow_cuts;
int onilda_flow_words(const int32_t* hour, const int32_t* minute, const int32_t* second, const int64_t* ibyt,
                      const int64_t* ipkt, const int32_t* sport, const int32_t* dport, int64_t n, int device,
                      onilda_flow_cuts* cuts_out, int32_t* src_code, int32_t* dst_code);
/* Columns and codes in device memory; cuts_out on the host (the call synchronises `stream`). */
int onilda_flow_words_device(const int32_t* hour, const int32_t* minute, const int32_t* second, const int64_t* ibyt,
                             const int64_t* ipkt, const int32_t* sport, const int32_t* dport, int64_t n,
                             onilda_flow_cuts* cuts_out, int32_t* src_code, int32_t* dst_code, void* stream);
/* Returns the string's length, or a negative error code. */
int onilda_flow_word_string(int32_t code, char* buf, size_t cap);

/* DNSWordCreation.dnsWordCreation (dns/DNSWordCreation.scala:32-145) on a byte buffer of query names
 * (names[name_offsets[i] .. name_offsets[i+1]), at most 255 bytes each, compared as bytes):
 * extractSubdomain (:155-196, String.split's dropping of trailing empty parts and the literal "None" of a
 * missing domain / subdomain included), Entropy.stringEntropy of the subdomain column
 * (utilities/Entropy.scala:16-23; "None" has entropy 2.0 and takes part in the cuts, as there),
 * getTopDomain (:147-153) against the caller's list of top-domain labels (the driver already reduces
 * top-1m.csv to them, :42-45), time and frame-length deciles, quintiles of subdomain length, entropy and
 * number of parts over their entries > 0 (:82-121), Quantiles.bin of all five and dnsWord (:226-235).
 * One 64-bit code per event:
 *   ((((((top*10 + frameBin)*10 + timeBin)*5 + subdomainLengthBin)*5 + entropyBin)*5 + periodsBin) << 32)
 *     | queryType << 16 | responseCode        (query type and response code 0..65535)
 * onilda_dns_word_string prints it as the reference's word ("1_4_7_0_2_3_1_0").  Entropy terms are added
 * largest count first (the reference adds in the unspecified iteration order of a Scala Map: the last bit
 * of an entropy may differ from a JVM run, the value as a function of the character counts does not). */
typedef struct onilda_dns_cuts {
    double time[10];
    double frame_length[10];
    double subdomain_length[5];
    double entropy[5];
    double num_periods[5];
} onilda_dns_cuts;
int onilda_dns_words(const char* names, const int64_t* name_offsets, const int32_t* frame_len,
                     const int64_t* unix_tstamp, const int32_t* qry_type, const int32_t* qry_rcode, int64_t n,
                     const char* top_domains, const int64_t* top_offsets, int64_t n_top, int device,
                     onilda_dns_cuts* cuts_out, int64_t* word_code);
/* The same, also returning the per-event columns DNSPostLDA keeps beside the word (any of them may be null):
 * subdomain length, number of parts, subdomain entropy (doubles, as the reference's columns), the top-domain
 * flag 0/1/2, and spans_out[n][3] = domain begin, domain end, subdomain end as byte positions inside the
 * name (the subdomain starts at 0; -1 where the reference has "None"). */
int onilda_dns_words_features(const char* names, const int64_t* name_offsets, const int32_t* frame_len,
                              const int64_t* unix_tstamp, const int32_t* qry_type, const int32_t* qry_rcode, int64_t n,
                              const char* top_domains, const int64_t* top_offsets, int64_t n_top, int device,
                              onilda_dns_cuts* cuts_out, int64_t* word_code, double* subdomain_length_out,
                              double* num_parts_out, double* entropy_out, int32_t* top_out, int32_t* spans_out);
/* Everything in device memory except cuts_out; top_slots: open-addressing table over FNV-1a 64 of the
 * labels (entry = label index or -1, top_capacity a power of two, 0 = no list), as onilda_dns_words builds. */
int onilda_dns_words_device(const char* names, const int64_t* name_offsets, const int32_t* frame_len,
                            const int64_t* unix_tstamp, const int32_t* qry_type, const int32_t* qry_rcode, int64_t n,
                            const char* top_bytes, const int64_t* top_offsets, const int32_t* top_slots,
                            int64_t top_capacity, onilda_dns_cuts* cuts_out, int64_t* word_code, void* stream);
int onilda_dns_word_string(int64_t code, char* buf, size_t cap);

/* Word codes are below this (the size of the code space above). */
#define ONILDA_FLOW_CODE_SPACE 65538000
/* FlowPreLDA.flowPreLDA (netflow/FlowPreLDA.scala:128-139) + the word dictionary and createModel of
 * OniLDACWrapper (OniLDACWrapper.scala:46-68, 146-168) in one call, with everything between the event
 * columns and the corpus kept on the device: flow words, the dictionary (word id = rank of the code
 * among the distinct codes, ascending; the codes come back in vocab_codes_out[0 .. *num_terms_out),
 * room for vocab_cap), word counts of the source role and of the destination role, their union grouped
 * into the corpus.  sip / dip are the caller's dense address ids (the ip dictionary stays with the
 * caller); num_docs 0 = max id + 1.  src_word_out / dst_word_out (n each, may be null) are the events'
 * word ids, which the scorer needs later.  n < 2^30. */
int onilda_flow_prelda(const int32_t* hour, const int32_t* minute, const int32_t* second, const int64_t* ibyt,
                       const int64_t* ipkt, const int32_t* sport, const int32_t* dport, const int32_t* sip,
                       const int32_t* dip, int64_t n, int64_t num_docs, int device, onilda_flow_cuts* cuts_out,
                       onilda_corpus** corpus_out, int32_t* vocab_codes_out, int32_t vocab_cap, int32_t* num_terms_out,
                       int32_t* src_word_out, int32_t* dst_word_out);

/* Word counts: replaces `.map((ip, word) -> 1).reduceByKey(_ + _)` of FlowPreLDA.scala:131-137 and
 * DNSPreLDA.scala:114-120 on dense integer ids (the string dictionaries stay with the caller,
 * OniLDACWrapper.scala:40-60).  n (document id, word id) occurrences -> the distinct pairs with their
 * number of occurrences, sorted by (document, word).  The *_out arrays need room for n entries;
 * *n_out = number of distinct pairs.  num_docs / num_terms 0 = max id + 1.  n < 2^31 - 1.
 * Flow calls it twice (source pairs, destination pairs: the reference reduces the two roles
 * separately and unions them, which is where a document's repeated word ids come from) and hands
 * the concatenation to onilda_corpus_from_triples. */
int onilda_pairs_count(const int32_t* doc, const int32_t* word, int64_t n, int64_t num_docs, int32_t num_terms,
                       int device, int32_t* doc_out, int32_t* word_out, int32_t* count_out, int64_t* n_out);
/* Same with device pointers on `stream`, no synchronisation; ids must be in range; *n_out is device memory. */
int onilda_pairs_count_device(const int32_t* doc, const int32_t* word, int64_t n, int64_t num_docs, int32_t num_terms,
                              int32_t* doc_out, int32_t* word_out, int32_t* count_out, long long* n_out, void* stream);
/* The same grouping with everything in device memory, ordered on `stream` (a cudaStream_t), no
 * synchronisation: offsets_out[num_docs + 1], words_out[n], counts_out[n]; totals_out[num_docs]
 * (int sum of counts per document, lda.h:27) and max_len_out[1] may be null.  Ids must already be in
 * range.  Scratch (about 16 bytes per triple) comes from the device's stream-ordered pool. */
int onilda_triples_group_device(const int32_t* doc, const int32_t* word, const int32_t* count, int64_t n,
                                int64_t num_docs, int64_t* offsets_out, int32_t* words_out, int32_t* counts_out,
                                int32_t* totals_out, int32_t* max_len_out, void* stream);
int onilda_corpus_dims(const onilda_corpus* c, int64_t* num_docs, int64_t* nnz, int32_t* num_terms,
                       int64_t* tokens, int32_t* max_len);
/* Borrowed views of the CSR arrays (valid until onilda_corpus_free). */
int onilda_corpus_csr(const onilda_corpus* c, const int64_t** offsets, const int32_t** words,
                      const int32_t** counts, const int32_t** totals);
/* Writes documents [doc_begin, doc_end) in the `.dat` form OniLDACWrapper.createModel
 * emits (OniLDACWrapper.scala:146-168): "<#entries> <word>:<count> ...\n" per document. */
int onilda_corpus_save_dat(const onilda_corpus* c, int64_t doc_begin, int64_t doc_end, const char* path);
/* Page-locks the corpus' host arrays (cudaHostRegister) so that uploads run at
 * full PCIe rate and asynchronously.  Optional; undone by onilda_corpus_free. */
int onilda_corpus_pin(onilda_corpus* c);
void onilda_corpus_free(onilda_corpus* c);

/* --------------------------------------------------------------------- model
 * replaces lda_model / new_lda_model / free_lda_model (lda.h:40-47,
 * lda-model.c:151-185).  new: alpha = 1.0, log_beta = 0. */
int onilda_model_new(int num_topics, int num_terms, onilda_model** out);
int onilda_model_dims(const onilda_model* m, int* num_topics, int* num_terms);
double onilda_model_alpha(const onilda_model* m);
int onilda_model_set_alpha(onilda_model* m, double alpha);
/* Borrowed host pointer to log_beta [K][V]. */
double* onilda_model_log_beta(onilda_model* m);
/* replaces seedMT + random_initialize_ss + lda_mle(.,.,0) (cokus.c:52-145,
 * lda-model.c:104-117, lda-estimate.c:155-162): same MT19937 stream, so a given
 * seed gives the reference's 000.beta.  alpha is left untouched (the caller
 * sets INITIAL_ALPHA). */
int onilda_model_init_random(onilda_model* m, uint32_t seed);
/* replaces corpus_initialize_ss + lda_mle (lda-model.c:120-144, lda-estimate.c:147-154). */
int onilda_model_init_seeded(onilda_model* m, const onilda_corpus* c, uint32_t seed);
/* replaces load_lda_model (lda-model.c:220-251): alpha and every beta value are
 * parsed as FLOAT and widened, as the reference does. */
int onilda_model_load(const char* model_root, onilda_model** out);
/* replaces save_lda_model (lda-model.c:193-217): <root>.beta (" %5.10f" per
 * value, one line per topic) and <root>.other. Byte-identical formatting. */
int onilda_model_save(const onilda_model* m, const char* model_root);
void onilda_model_free(onilda_model* m);
/* replaces save_gamma (lda-estimate.c:88-104): "%5.10f" then " %5.10f". */
int onilda_gamma_save(const char* path, const double* gamma, int64_t num_docs, int num_topics);

/* ------------------------------------------------------------------------ EM
 * One handle = run_em's state for the documents [doc_begin, doc_end) of `c` on
 * one device (lda-estimate.c:112-321).  The reference shards round-robin over
 * MPI ranks (:201); here a shard is a contiguous document range, one per GPU.
 * device < 0 selects the current device.  The start model is copied. */
int onilda_em_create(const onilda_corpus* c, int64_t doc_begin, int64_t doc_end, const onilda_model* start,
                     const onilda_settings* settings, int device, onilda_em** out);
void onilda_em_destroy(onilda_em* em);

/* E-step over the shard: lda_inference + compute_likelihood + doc_e_step for
 * every document (lda-inference.c:27-127, lda-estimate.c:28-60, 198-209), in
 * sm_100a kernels.  Leaves the shard's sufficient statistics on the device. */
int onilda_em_estep(onilda_em* em);
/* The shard's sufficient statistics as ONE packed fp64 device buffer, the
 * operand of the per-iteration sum over shards (replaces the 2K+3 MPI_Allreduce
 * calls at lda-estimate.c:211-252):
 *    [ class_word, word-major, (V + 16) x kpad (the last 16 rows zero) | likelihood | alpha_suffstats | num_docs | 0 ]
 * Single shard: skip this.  Several shards WITHOUT a communicator attached (see below): all-reduce (SUM)
 * `count` doubles in place across the shards' buffers (onilda_em_allreduce_peer, or the caller's own
 * collective), then call onilda_em_mstep on every shard. */
int onilda_em_suffstats_device(onilda_em* em, double** dev_ptr, int64_t* count);
/* M-step + convergence bookkeeping: lda_mle (lda-model.c:27-54, log-normalise
 * on device; opt_alpha's scalar Newton on the host, lda-alpha.c:42-68), then
 * converged = (old - new)/old, VAR_MAX_ITER doubling when negative, and the
 * loop test of lda-estimate.c:186 / 260-262.
 *   likelihood, converged : the values the reference prints to likelihood.dat
 *   keep_going            : 1 while run_em's while-condition still holds */
int onilda_em_mstep(onilda_em* em, double* likelihood, double* converged, int* keep_going);
/* Re-sends the shard's CSR arrays (from the corpus given to onilda_em_create,
 * which must still be alive) and `m` (alpha + log_beta) host -> device, and
 * resets the EM loop state: the handle is then as freshly created, without
 * reallocating.  Returns once the copies are done (the caller may reuse `m`).  With it one call sequence reload -> iterate -> get_model /
 * get_gamma is a stateless EM step on HOST buffers. */
int onilda_em_reload(onilda_em* em, const onilda_model* m);
/* The two halves of onilda_em_reload.  A caller that feeds a stream of corpora from host memory keeps two handles and
 * sends the next one's CSR arrays (reload_corpus, from another thread: handles are independent and each has its own
 * stream) while the current one iterates; the model, which depends on the step before, follows with reload_model.
 * No reference counterpart (the reference reads its corpus once, lda-estimate.c:434). */
int onilda_em_reload_corpus(onilda_em* em);
int onilda_em_reload_model(onilda_em* em, const onilda_model* m);
/* estep + mstep for a single shard. */
int onilda_em_iterate(onilda_em* em, double* likelihood, double* converged, int* keep_going);
/* In-process sum of the packed statistics of `n` handles (one per device, same
 * process) over NVLink peer memory; each device reduces its slice of the buffer
 * from all peers, always in rank order, and writes the sum back to every peer:
 * the alternative to the communicator below when identical bits from run to run
 * matter more than the replicated M-step it leaves to every device. */
int onilda_em_allreduce_peer(onilda_em** ems, int n);

/* Several shards, one communicator (NCCL over NVLink / NVSwitch).  Replaces the MPI calls of
 * lda-estimate.c:211-252 - MPI_Barrier, then MPI_Allreduce of alpha_suffstats, num_docs, K rows of
 * class_word, K class_total values and the likelihood - and the lda_mle every rank then repeats (:256).
 * With a communicator attached, onilda_em_mstep / onilda_em_mstep_group run on the handles' streams, with no
 * host synchronisation before the final read-back of four scalars:
 *     reduce-scatter of class_word (rank r receives the sums of rows [r*ceil(V/P), (r+1)*ceil(V/P)))
 *     -> class_total partial sums of those rows -> all-reduce of [likelihood, alpha_suffstats, num_docs, class_total]
 *     -> lda_mle of the own rows -> all-gather of log_beta, the E-step's beta' table and its row maxima.
 * The caller does NOT all-reduce onilda_em_suffstats_device itself in this mode.
 *   one process per GPU (torchrun, mpiexec): rank 0 calls onilda_comm_unique_id, ships the 128 bytes to the
 *     others by its own means, every rank calls onilda_em_comm_attach (collective, like ncclCommInitRank);
 *   one process, n devices: onilda_em_comm_attach_local on the n handles, then onilda_em_mstep_group on all of them.
 * At most 16 ranks.  Every rank must hold the same V and K. */
#define ONILDA_COMM_ID_BYTES 128
int onilda_comm_unique_id(unsigned char* id /* [ONILDA_COMM_ID_BYTES] */);
int onilda_em_comm_attach(onilda_em* em, const unsigned char* id, int nranks, int rank);
int onilda_em_comm_attach_local(onilda_em** ems, int n);
/* onilda_em_mstep for all handles of this process at once (NCCL calls of several devices issued by one thread
 * must sit in one group).  likelihood / converged / keep_going are those of the whole corpus. */
int onilda_em_mstep_group(onilda_em** ems, int n, double* likelihood, double* converged, int* keep_going);
/* What the last mstep's exchange took on this handle's stream, in ms (CUDA events; includes the wait for the
 * slowest rank's E-step), and the communicator's shape. */
int onilda_em_comm_stats(onilda_em* em, int* nranks, int* rank, double* reduce_scatter_ms, double* all_gather_ms);
/* The exchange by itself, for bus-bandwidth figures: `repeats` reduce-scatters of the statistics table, then as
 * many rounds of the three all-gathers, each set started together on all ranks; per-call ms (max over the
 * handles given) and the table bytes one call reduces / gathers.  Collective.  Leaves the statistics buffer
 * undefined until the next onilda_em_estep; the model is unchanged. */
int onilda_em_collective_time(onilda_em** ems, int n, int repeats, double* rs_ms, double* ag_ms, int64_t* bytes);

/* Current model (after the last mstep) and the variational gamma of the
 * shard's documents from the last estep ([doc_end-doc_begin][K], row i is
 * document doc_begin+i: line i of final.gamma). */
int onilda_em_get_model(onilda_em* em, onilda_model* out);
int onilda_em_get_gamma(onilda_em* em, double* gamma_out);
/* Per-document likelihood bound from the last estep ([doc_end-doc_begin]). */
int onilda_em_get_doc_likelihood(onilda_em* em, double* out);
/* Variational sweeps each document took in the last estep (lda_inference's var_iter,
 * lda-inference.c:48): [doc_end-doc_begin]. */
int onilda_em_get_doc_sweeps(onilda_em* em, int32_t* out);
/* How many (document, word) steps of the last estep were redone in log space because their sum over topics left
 * the range of the product form (a diagnostic; the reference computes every word that way, lda-inference.c:52-68). */
int onilda_em_log_space_words(onilda_em* em, int64_t* count);
/* Work of the last iteration's E-step: sum over all documents (of all shards, after onilda_em_mstep) of
 * entries x variational sweeps, i.e. how many times lda_inference's word loop body ran (lda-inference.c:50-75). */
int onilda_em_word_steps(onilda_em* em, double* steps);
/* Counters of the last estep: kernel time in ms (CUDA events on the handle's
 * stream), number of kernel launches issued by the last estep+mstep pair, and
 * total variational sweeps summed over documents. */
int onilda_em_stats(onilda_em* em, double* estep_ms, double* mstep_ms, int64_t* launches, int64_t* sweeps);
/* The handle's stream as a cudaStream_t (for callers that enqueue their own
 * collective on it). */
void* onilda_em_stream(onilda_em* em);

/* run_em end to end on `ndev` devices of this process (lda-estimate.c:112-321):
 * writes <dir>/000.{beta,other}, likelihood.dat ("%10.10f\t%5.5e\n" per
 * iteration), final.{beta,other,gamma}.  start is "random", "seeded" or a model
 * root, as argv[7] of `lda est`.  seed feeds the MT19937 of the two inits. */
int onilda_run_em(const char* start, const char* directory, const onilda_corpus* c, int num_topics,
                  double initial_alpha, const onilda_settings* settings, int ndev, uint32_t seed);

/* ----------------------------------------------------------------- inference
 * replaces infer (lda-estimate.c:358-388): per-document lda_inference with a
 * fixed model.  gamma_out [D][K], likelihood_out [D]. */
int onilda_infer(const onilda_model* m, const onilda_corpus* c, const onilda_settings* settings, int device,
                 double* gamma_out, double* likelihood_out);
/* The file-writing form `lda inf` uses: <name>-lda-lhood.dat ("%5.5f\n") and
 * <name>-gamma.dat. */
int onilda_infer_files(const char* model_root, const char* name, const onilda_corpus* c,
                       const onilda_settings* settings, int device);

/* ------------------------------------------------------------------- scoring
 * replaces OniLDACWrapper.getDocumentResults / getTopicDocument
 * (OniLDACWrapper.scala:132-144, 170-176): theta = gamma / sum, zeros if sum <= 0. */
int onilda_theta_from_gamma(const double* gamma, int64_t num_docs, int num_topics, int device, double* theta_out);
/* replaces getWordProbabilitesFromTopicLine + getWordToProbPerTopicMap
 * (OniLDACWrapper.scala:121-130, 178-195): phi[w][k] = exp(lb[k][w]) / sum_w exp(lb[k][w]). */
int onilda_phi_from_log_beta(const double* log_beta, int num_topics, int num_terms, int device, double* phi_out);

/* replaces the broadcast maps of SuspiciousConnectsScoreFunction
 * (SuspiciousConnectsScoreFunction.scala:7-9): tables resident on one device. */
int onilda_scorer_create(const double* theta, int64_t num_docs, const double* phi, int32_t num_terms,
                         int num_topics, int device, onilda_scorer** out);
/* The same tables taken straight from an EM handle's device memory (the handle's shard: its gamma rows
 * become theta rows 0..D-1, its current model becomes phi): what getDocumentResults +
 * getWordToProbPerTopicMap compute after reading final.gamma / final.beta back
 * (OniLDACWrapper.scala:88-144, 170-195), without the text files or a host round trip. */
int onilda_scorer_create_from_em(onilda_em* em, onilda_scorer** out);
void onilda_scorer_destroy(onilda_scorer* s);
/* replaces SuspiciousConnectsScoreFunction.score (:11-23) for n (doc, word)
 * pairs: sum_k theta[doc][k] * phi[word][k], products rounded then summed left
 * to right as the Scala zip/map/sum does.  An id outside the table (e.g. -1)
 * takes the uniform 1/K row.  Host buffers; copies are pipelined inside. */
int onilda_score(onilda_scorer* s, const int32_t* doc, const int32_t* word, int64_t n, double* out);
/* replaces FlowPostLDA.score x2 + minimumScore (netflow/FlowPostLDA.scala:94-125):
 * min(score(sdoc, sword), score(ddoc, dword)) per flow event.
 * Unknown ids: a word id outside the table scores against the uniform row (:100, Array.fill(20)(0.05)).
 * An ip (document id) outside the table is different for flow: the reference inner-joins the events with the
 * document table (:94), once per side, so such an event is absent from its results.  out[] keeps one slot per
 * event, so that event's slot holds +infinity; onilda_topk / onilda_flow_results / onilda_score_flow_topk_device
 * never return a +infinity row, whatever the threshold.  (onilda_score - dns, proxy - gives an unknown ip the
 * uniform row, as SuspiciousConnectsScoreFunction.scala:17 does.) */
int onilda_score_flow_min(onilda_scorer* s, const int32_t* sdoc, const int32_t* sword, const int32_t* ddoc,
                          const int32_t* dword, int64_t n, double* out);
/* replaces filter(score <= threshold) + orderBy/takeOrdered(k)
 * (netflow/FlowPostLDA.scala:79-80, dns/DNSPostLDA.scala:49-70,
 * utilities/DataFrameUtils.scala:21-38): ascending, ties by event index,
 * k == -1 keeps all.  idx_out/score_out need room for min(k, n) (n when k == -1). */
int onilda_topk(onilda_scorer* s, const double* score, int64_t n, double threshold, int64_t k,
                int64_t* idx_out, double* score_out, int64_t* n_out);
/* SURVEY section 8(f) row N3.  FlowPostLDA.scala:75-80 in one call: SourceScore, DestinationScore,
 * their minimum, filter(min <= threshold), orderBy(min), limit(k).  Per kept row, ascending by the
 * minimum (ties by event index): the event's index (to join its own columns), its source score and
 * its destination score - the two score columns of flow_results.csv (the minimum is dropped from
 * the file, :80).  Outputs need room for min(k, n) rows (n when k == -1). */
int onilda_flow_results(onilda_scorer* s, const int32_t* sdoc, const int32_t* sword, const int32_t* ddoc,
                        const int32_t* dword, int64_t n, double threshold, int64_t k, int64_t* idx_out,
                        double* src_score_out, double* dst_score_out, int64_t* n_out);
/* The rows of dns_results.csv / proxy_results.csv in one call: score per event (one side: client ip x word), keep
 * score <= threshold, the k smallest in ascending order; k == -1 keeps all (replaces DNSPostLDA.scala:44-70 - filter,
 * count, takeOrdered(takeCount), sortBy - ProxySuspiciousConnectsAnalysis.scala:45-51 and
 * DataFrameUtils.dfTakeOrdered, utilities/DataFrameUtils.scala:21-38).  idx_out: event indices of the kept rows (join
 * the event's own columns on the caller's side), score_out: the Score column.  Equal scores come out by event
 * index (Spark leaves them in no particular order).  Outputs need room for min(k, n) entries, n when k == -1. */
int onilda_results_rows(onilda_scorer* s, const int32_t* doc, const int32_t* word, int64_t n, double threshold,
                        int64_t k, int64_t* idx_out, double* score_out, int64_t* n_out);
/* Writes <type>_results.csv: line i = rows[i] + delimiter + score_a[i] [+ delimiter + score_b[i]] (replaces the
 * `_.mkString(outputDelimiter)` + saveAsTextFile of netflow/FlowPostLDA.scala:80, dns/DNSPostLDA.scala:69,
 * proxy/ProxySuspiciousConnectsAnalysis.scala:51; the part files ml_ops.sh merges are one file here).
 *   rows[i]  : the kept event's own columns and word column(s), already joined with the delimiter by the caller,
 *              in the order onilda_flow_results / onilda_results_rows returned them;
 *   flow     : score_a = SourceScore, score_b = DestinationScore (columns 35 / 36 for oni-oa, oa/flow/flow_conf.json);
 *   dns, proxy: score_a = Score, score_b = NULL.
 * Doubles are printed in Java's Double.toString layout (see onilda_format_java_double).  Temporary file + rename. */
int onilda_results_write(const char* path, const char* const* rows, int64_t m, const double* score_a, const double* score_b,
                         const char* delimiter);
/* v as Java's Double.toString lays it out, with the shortest digits that read back as v ("0.05", "1.0E-4",
 * "6.379973716629045E-4", "1.2345678E7").  out: at least 32 bytes; returns the length, no NUL.  JDK 19+ prints
 * exactly this; the JDK 7/8 that Spark 1.6 ran on prints one more digit for a small share of values. */
int onilda_format_java_double(double v, char* out);
/* Fused form for data already on the device (no host copies): scores the
 * events, keeps the running top-k, returns only the k rows.  Device pointers. */
int onilda_score_flow_topk_device(onilda_scorer* s, const int32_t* sdoc_dev, const int32_t* sword_dev,
                                  const int32_t* ddoc_dev, const int32_t* dword_dev, int64_t n,
                                  double threshold, int64_t k, int64_t* idx_out, double* score_out,
                                  int64_t* n_out, double* score_dev_out /* may be NULL */);
int onilda_scorer_stats(onilda_scorer* s, double* kernel_ms, int64_t* launches);

/* Top-k for callers that cannot allocate.  (No counterpart in the reference: it exists because
 * of where this library runs, not because of anything lda-c or oni-ml does.)
 *
 * Who that is: code running inside a process that has already given all of the GPU's memory to
 * a pooling allocator - a Spark executor with a GPU memory pool, which is where ONI's post-LDA
 * scoring lives - so that a cudaMalloc made by this library would fail; and latency-bound
 * streaming scorers, because cudaMalloc / cudaFree stall the whole device.
 * onilda_topk and onilda_score_flow_topk_device both allocate.
 *
 * This one makes no cudaMalloc / cudaFree / host-heap call.  Same result as onilda_topk (the k
 * smallest scores that are <= threshold, ascending, ties by event index), with
 *   - score_dev: DEVICE pointer to n scores on the current device (n < 2^44); idx_out,
 *     score_out: HOST arrays with room for k entries; *n_out receives the number written;
 *   - 0 <= k <= ONILDA_NOALLOC_MAX_K; a larger k, or -1, is refused with ONILDA_EINVAL and a
 *     message naming the limit;
 *   - its scratch (histogram, counter, the selected rows: 82 KB) is a fixed device array inside
 *     the library, one per GPU, kept until the process exits.  Every call on a GPU shares it and
 *     finds the previous call's contents there; the entry point clears what has to start at
 *     zero itself.  Two calls that overlap on the same GPU (two host threads, or one thread and
 *     work of its own still running that reads the results) race on that array: serialise them.
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress);
 *     make that first call before starting any stream capture.
 *   - the scores must be COMPLETE before the call: synchronise the stream that produced them.
 *     The work here goes to the legacy default stream, which does not wait for streams created
 *     with cudaStreamNonBlocking - the scorer's own streams are of that kind.
 *   - the call is synchronous: it reads small results back with blocking copies between its
 *     passes and returns when the rows are in the host arrays.
 *   - about 80 KB of the CALLER'S STACK are used (one histogram and the k rows): mind threads
 *     with small stacks, e.g. JVM threads that come in through JNI. */
#define ONILDA_NOALLOC_MAX_K 4096
int onilda_topk_device_noalloc(const double* score_dev, int64_t n, double threshold, int64_t k,
                               int64_t* idx_out, double* score_out, int64_t* n_out);

#ifdef __cplusplus
}
#endif
#endif /* ONILDA_H */
