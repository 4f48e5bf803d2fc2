/*
 * guac_b200.h — C ABI of libguac_b200.so
 *
 * B200-native (sm_100a) implementation of Guacamole's per-locus hot path:
 * pileup construction -> genotype / somatic likelihoods.  Every entry point
 * names the reference (hammerlab/guacamole, Scala) interface it replaces as
 * `file:line` relative to src/main/scala/org/hammerlab/guacamole/ ("M/").
 *
 * Conventions
 *   - every function returns GUAC_OK (0) or a negative guac_status; the
 *     message is available from guac_last_error(ctx).  No C++ exception
 *     crosses this boundary.
 *   - inputs are caller-owned host memory and must stay valid until the call
 *     returns; calls are synchronous at return.
 *   - outputs are library-owned host memory, released with guac_free().
 *   - a guac_ctx is bound to one CUDA device and is not thread-safe.
 *   - loci are 0-based, ranges are half-open [start, end), exactly like
 *     the reference's Locus / LociSet.
 *   - there is no CPU fallback: without a CUDA device guac_ctx_create fails.
 */
#ifndef GUAC_B200_H
#define GUAC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct guac_ctx guac_ctx;

typedef enum {
  GUAC_OK = 0,
  GUAC_E_INVALID_INPUT = -1, /* reference-side assume()/assert()/InvalidCigarElementException */
  GUAC_E_CUDA = -2,
  GUAC_E_NO_DEVICE = -3,
  GUAC_E_OOM = -4,
  GUAC_E_UNSUPPORTED = -5, /* a documented capacity limit of this library was exceeded */
  GUAC_E_STATE = -6        /* e.g. no reads / reference uploaded yet */
} guac_status;

/* CIGAR operator codes: BAM encoding, word = (length << 4) | op.
 * (htsjdk CigarOperator, used by M/pileup/PileupElement.scala:59-117) */
enum { GUAC_CIGAR_M = 0, GUAC_CIGAR_I = 1, GUAC_CIGAR_D = 2, GUAC_CIGAR_N = 3, GUAC_CIGAR_S = 4,
       GUAC_CIGAR_H = 5, GUAC_CIGAR_P = 6, GUAC_CIGAR_EQ = 7, GUAC_CIGAR_X = 8 };

/* per-read flag bits (M/reads/MappedRead.scala:21-33) */
enum { GUAC_READ_POSITIVE_STRAND = 1, GUAC_READ_PAIRED = 2, GUAC_READ_DUPLICATE = 4, GUAC_READ_FAILED_QC = 8 };

/*
 * Columnar (struct-of-arrays) batch of mapped reads: the data contract of
 * M/reads/MappedRead.scala:21-33 for one partition of
 * M/readsets/rdd/PartitionedRegions.scala:56 (reads of all samples, sorted by
 * (contig_id, start); every read that overlaps the partition's loci is present,
 * halo reads included).
 *
 * Bases are 2 bit/base (A=0 C=1 G=2 T=3), base i of the batch in bits
 * [2*(i%4), 2*(i%4)+1] of bases2[i/4].  Bases other than upper-case ACGT are
 * stored as 0 in bases2, have their bit set in nmask (bit i%8 of nmask[i/8]) and
 * are listed in (exc_pos, exc_base), sorted by exc_pos (batch-global base index).
 * guac_pack_bases() produces this encoding from ASCII.
 */
typedef struct {
  int64_t n_reads;
  int32_t n_samples;
  const int32_t* contig_id;  /* [n_reads] */
  const int64_t* start;      /* [n_reads] 0-based alignment start */
  const uint32_t* cigar_off; /* [n_reads+1] offsets into cigar */
  const uint32_t* cigar;     /* BAM-encoded operators */
  const uint64_t* seq_off;   /* [n_reads+1] offsets into the base/qual streams, in bases */
  const uint8_t* bases2;     /* [(seq_off[n]+3)/4] */
  const uint8_t* nmask;      /* [(seq_off[n]+7)/8] */
  int64_t n_exc;
  const uint64_t* exc_pos;   /* [n_exc] */
  const uint8_t* exc_base;   /* [n_exc] ASCII */
  const uint8_t* qual;       /* [seq_off[n]] Phred, not +33 */
  const uint8_t* mapq;       /* [n_reads] */
  const uint8_t* flags;      /* [n_reads] GUAC_READ_* */
  const uint8_t* sample_id;  /* [n_reads] < n_samples */
} guac_read_batch;

/* One contiguous run of loci, the unit of M/loci/partitioning/LociPartitioning.scala:25-40 */
typedef struct {
  int32_t contig_id;
  int32_t _pad;
  int64_t start;
  int64_t end;
} guac_range;

/* ---- context ------------------------------------------------------------ */

/* Binds a context to CUDA device `device`.  Fails with GUAC_E_NO_DEVICE when no
 * GPU is present (no CPU fallback).  Replaces the per-task setup of
 * M/readsets/rdd/PartitionedRegions.scala:56-72. */
int guac_ctx_create(int device, guac_ctx** out);
void guac_ctx_destroy(guac_ctx* ctx);
const char* guac_last_error(guac_ctx* ctx);
/* Releases any buffer handed out through an out-pointer of this API. */
void guac_free(guac_ctx* ctx, void* p);
/* Number of kernels this context launched so far (bench.py's gpu_launches). */
int64_t guac_kernel_launches(guac_ctx* ctx);

/* ---- host-side packing helper (no device work) ---------------------------
 * ASCII bases -> (bases2, nmask, exceptions).  `base_index0` is the batch-global
 * index of ascii[0] and must be a multiple of 8.  exc_pos/exc_base must have room
 * for exc_cap entries; *n_exc receives the number found (GUAC_E_OOM if > cap).
 * Mirrors what M/reads/Read.scala:63-116 does when it materialises `sequence`. */
int guac_pack_bases(const uint8_t* ascii, int64_t n, uint64_t base_index0, uint8_t* bases2, uint8_t* nmask,
                    uint64_t* exc_pos, uint8_t* exc_base, int64_t exc_cap, int64_t* n_exc);

/* ---- reference ------------------------------------------------------------
 * M/reference/ReferenceBroadcast.scala:42-83 (ContigSequence.apply / slice).
 * guac_ref_reserve declares the window [offset, offset+len) of a contig that will
 * be addressable; guac_ref_upload copies upper-case ASCII bases into it.  Loci of
 * the window that were never uploaded read as "missing" and make any pileup that
 * touches them fail with GUAC_E_INVALID_INPUT, like MapBackedReferenceSequence. */
int guac_ref_reserve(guac_ctx* ctx, int32_t contig_id, int64_t offset, int64_t len);
int guac_ref_upload(guac_ctx* ctx, int32_t contig_id, const uint8_t* bases_ascii_upper, int64_t len,
                    int64_t offset);

/* ---- reads ---------------------------------------------------------------- */
/* Copies one partition's reads to the device (async H2D on the context's stream)
 * and derives per-read end loci.  Replaces the Iterator[MappedRead] handed to the
 * closure of M/readsets/rdd/PartitionedRegions.scala:56-72. */
int guac_reads_upload(guac_ctx* ctx, const guac_read_batch* batch);

/* ---- pileup --------------------------------------------------------------- */

/* Per visited locus, per sample summary of the pileup
 * (M/pileup/Pileup.scala:32-52,100-106 — depth, positiveDepth, referenceDepth — and the
 * alignment classes of M/pileup/Alignment.scala:14-75). */
typedef struct {
  uint32_t depth;          /* Pileup.depth */
  uint32_t reference_depth;/* Pileup.referenceDepth: Match elements only */
  uint32_t positive_depth; /* Pileup.positiveDepth */
  uint32_t mismatches;     /* Mismatch */
  uint32_t insertions;     /* Insertion */
  uint32_t deletions;      /* Deletion (anchor base before the deleted run) */
  uint32_t mid_deletions;  /* MidDeletion */
  uint32_t clipped;        /* Clipped (N skips) */
} guac_locus_counts;

/*
 * pileupFlatMap{OneSample,TwoSamples,MultipleSamples} with a counting closure
 * (M/distributed/PileupFlatMapUtils.scala:52,79,105; locus visitation of
 * M/windowing/SlidingWindow.scala:124-162 with halfWindowSize = 0).
 * Visits, in order, every locus of `loci` (sorted, disjoint) — with skip_empty != 0
 * only those where at least one sample has a read overlapping the locus.
 * Outputs: *n_out loci; out_contig[i], out_locus[i]; out_counts[i*n_samples + s].
 */
int guac_pileup_counts(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, int32_t skip_empty,
                       int32_t** out_contig, int64_t** out_locus, guac_locus_counts** out_counts, int64_t* n_out);

/* One PileupElement (M/pileup/PileupElement.scala:24-31) plus its classification. */
typedef struct {
  int64_t read_index;                 /* index into the uploaded batch */
  int32_t read_position;              /* readPosition */
  int32_t cigar_element_index;        /* cigarElementIndex */
  int64_t cigar_element_locus;        /* cigarElementLocus */
  int32_t index_within_cigar_element; /* indexWithinCigarElement */
  uint8_t kind;                       /* guac_alignment_kind */
  uint8_t base;                       /* Match/Mismatch: sequenced base (ASCII); else 0 */
  uint8_t quality;                    /* PileupElement.qualityScore :149-154 */
  uint8_t sample_id;
  int32_t indel_length;               /* Insertion: inserted bases; Deletion: deleted bases; else 0 */
  int32_t _pad;
} guac_element;

typedef enum {
  GUAC_ALN_MATCH = 0, GUAC_ALN_MISMATCH = 1, GUAC_ALN_INSERTION = 2, GUAC_ALN_DELETION = 3,
  GUAC_ALN_MID_DELETION = 4, GUAC_ALN_CLIPPED = 5
} guac_alignment_kind;

/*
 * Same operator, materialising every PileupElement so that a host closure can be
 * run on real Pileup objects (generic-closure path of PileupFlatMapUtils).
 * Elements of locus i, all samples, ordered by read_index, are
 * out_elements[out_offsets[i] .. out_offsets[i+1]).
 */
int guac_pileup_elements(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, int32_t skip_empty,
                         int32_t** out_contig, int64_t** out_locus, int64_t** out_offsets,
                         guac_element** out_elements, int64_t* n_out);

/*
 * The counting closure for hosts that must not allocate.  (No counterpart in the reference: the JVM has no
 * such constraint.)  Some callers cannot call cudaMalloc when they need a pileup — a plugin inside an
 * executor whose device memory belongs entirely to the host framework's pool allocator — so this variant
 * of guac_pileup_counts (dense: skip_empty = 0; ONE range of ONE contig) takes no guac_ctx and allocates
 * nothing, neither on the device nor pinned on the host: its working memory is a fixed array inside the
 * library, about 3.9 MiB per GPU that loads it, kept until the process exits.
 *
 *   d_batch      a guac_read_batch whose POINTERS ARE DEVICE POINTERS (the struct itself is on the host);
 *                every read lies on `contig_id`, sorted by start
 *   d_ref_bases  device pointer to upper-case ASCII reference bases of loci [ref_offset, ref_offset+ref_len)
 *   d_out_counts device buffer of (range_end - range_start) * n_samples records, written in locus order
 *   stream       the cudaStream_t to run on (NULL = default stream); the call waits on it twice and returns
 *                after it has drained
 * Buffers need exactly the sizes guac_read_batch states (no padding).  The call makes `device` current while
 * it runs and restores the caller's current device before it returns.
 *
 * Limits: at most guac_noalloc_max_reads() reads (131 072) and guac_noalloc_max_loci() loci (524 288) per
 * call; larger inputs are refused with GUAC_E_UNSUPPORTED and a message naming the limit; negative reference
 * offsets / lengths and loci at or beyond 2^40 are GUAC_E_INVALID_INPUT.  The first call on a device also
 * resolves the array's device address; make that call before starting any stream capture.
 * All calls on a device share the array and it keeps the last call's contents (the call clears what must
 * start at zero, including d_out_counts), so calls whose GPU work overlaps — two streams, two threads —
 * race; serialise them.  Errors: status code, text from guac_noalloc_last_error().
 */
int guac_pileup_counts_noalloc(int device, const guac_read_batch* d_batch, const uint8_t* d_ref_bases, int64_t ref_offset,
                               int64_t ref_len, int32_t contig_id, int64_t range_start, int64_t range_end,
                               guac_locus_counts* d_out_counts, void* stream);
const char* guac_noalloc_last_error(void);
int64_t guac_noalloc_max_reads(void);
int64_t guac_noalloc_max_loci(void);

/* ---- callers -------------------------------------------------------------- */

/* M/variants/AlleleEvidence.scala:25-35 */
typedef struct {
  double probability;
  int32_t read_depth;
  int32_t allele_read_depth;
  int32_t forward_depth;
  int32_t allele_forward_depth;
  double mean_mapping_quality;
  double median_mapping_quality;
  double mean_base_quality;
  double median_base_quality;
  double median_mismatches_per_read;
} guac_allele_evidence;

/* SomaticStandard.Arguments (M/commands/SomaticStandardCaller.scala:35-73) */
typedef struct {
  int32_t normal_sample_id;                  /* sample_id of the normal reads (pileupFlatMapTwoSamples sample1) */
  int32_t tumor_sample_id;
  int32_t normal_odds_threshold;             /* --normal-odds, default 5 */
  int32_t tumor_odds_threshold;              /* --tumor-odds, default 10 */
  int32_t max_read_depth;                    /* maxTumorReadDepth, default INT32_MAX */
  int32_t max_normal_alternate_read_depth;   /* default 4 */
  float min_tumor_variant_allele_frequency;  /* percent / 100.0f, default 0.03f */
  int32_t _pad;
} guac_ss_params;

/* M/variants/CalledSomaticAllele.scala:25-34.  ref/alt are (offset,len) into the
 * returned string arena. */
typedef struct {
  int32_t contig_id;
  int32_t _pad;
  int64_t start;
  uint32_t ref_off, ref_len, alt_off, alt_len;
  double somatic_log_odds;
  double tumor_ref_log_prob, tumor_alt_log_prob, normal_ref_log_prob, normal_alt_log_prob;
  guac_allele_evidence tumor_variant_evidence;
  guac_allele_evidence normal_reference_evidence;
} guac_ss_call;

/* Dense per-(visited locus, sample) summary written by the fused callers. */
typedef struct {
  uint16_t depth;           /* saturating at 65535 */
  uint16_t reference_depth;
  uint16_t top_alt_count;   /* count of the most frequent variant allele */
  uint16_t flags;           /* bit0: likelihoods were evaluated here; bit1: a call was emitted */
} guac_locus_summary;

/*
 * somatic-standard: pileupFlatMapTwoSamples(normal, tumor, skipEmpty = true,
 * findPotentialVariantAtLocus) — M/commands/SomaticStandardCaller.scala:100-117,155-253,
 * with M/likelihood/Likelihood.scala:99-210 and M/variants/AlleleEvidence.scala:42-85
 * fused on the device.  Calls come back sorted by (range order, locus).
 * out_summary (optional, may be NULL) receives [*n_visited * 2] records, sample order
 * (normal, tumor); out_n_visited (optional) the number of loci visited.
 */
int guac_somatic_standard(guac_ctx* ctx, const guac_ss_params* params, const guac_range* loci, int32_t n_ranges,
                          guac_ss_call** out_calls, int64_t* n_calls, char** out_arena, int64_t* arena_len,
                          guac_locus_summary** out_summary, int64_t* out_n_visited);

/* ---- germline toy caller ------------------------------------------------------
 * The "germline-standard" path that survives in the reference: the shortcut branch of
 * GermlineAssemblyCaller.discoverGermlineVariants (M/commands/GermlineAssemblyCaller.scala:103-153). */
typedef struct {
  int32_t sample_id;
  int32_t min_alt_reads;               /* minAltReads, default 2 (:92) */
  int32_t min_phred_scaled_likelihood; /* --min-likelihood, default 0 (:46) */
  int32_t respect_active_regions;      /* 1: loci whose region isActiveRegion (M/assembly/AssemblyUtils.scala:28-39) are
                                          returned uncalled for the assembler, as with --shortcut-assembly; 0: call them too */
  float min_area_vaf;                  /* --min-area-vaf percent / 100.0f (:70), default 0.05f */
  int32_t _pad;
} guac_gs_params;

/* M/variants/CalledAllele.scala:24-31 */
typedef struct {
  int32_t contig_id;
  int32_t n_genotypes;             /* genotypes enumerated at the locus */
  int64_t start;
  uint32_t ref_off, ref_len, alt_off, alt_len;
  double genotype_log_probability; /* normalised natural-log probability of the called genotype */
  guac_allele_evidence evidence;
} guac_called_allele;

/*
 * windowFlatMapWithState over one sample with the closure: if depth - referenceDepth >= minAltReads then
 * callPileupVariant (M/commands/GermlineAssemblyCaller.scala:135-153,233-261), i.e.
 * Likelihood.probabilitiesOfAllPossibleGenotypesFromPileup (M/likelihood/Likelihood.scala:31-63), the most
 * probable genotype's non-reference alleles with their AlleleEvidence, kept when
 * phredScaledLikelihood > min_phred_scaled_likelihood.  out_active_loci (contig, locus pairs) lists the loci
 * left to the assembler when respect_active_regions is set.
 */
int guac_germline_standard(guac_ctx* ctx, const guac_gs_params* params, const guac_range* loci, int32_t n_ranges,
                           guac_called_allele** out_calls, int64_t* n_calls, char** out_arena, int64_t* arena_len,
                           int64_t** out_active_loci, int64_t* n_active, int64_t* out_n_visited);

/* ---- germline-assembly ---------------------------------------------------------
 * M/assembly/AssemblyArgs.scala:13-45 and GermlineAssemblyCaller.Arguments (M/commands/GermlineAssemblyCaller.scala:33-47). */
typedef struct {
  int32_t sample_id;
  int32_t kmer_size;                   /* --kmer-size, default 45 */
  int32_t assembly_window_range;       /* --assembly-window-range (= halfWindowSize), default 20; must be >= kmer_size */
  int32_t min_occurrence;              /* --min-occurrence, default 3 */
  int32_t min_mean_kmer_quality;       /* --min-mean-kmer-quality, default 0 */
  int32_t min_alt_reads;               /* 2 */
  int32_t min_phred_scaled_likelihood; /* --min-likelihood, default 0 */
  int32_t shortcut_assembly;           /* --shortcut-assembly */
  float min_area_vaf;                  /* --min-area-vaf percent / 100.0f */
  int32_t _pad;
} guac_ga_params;

/* One assembled window (the half-open reference slice [locus - range, locus + range) around a candidate locus). */
typedef struct {
  int32_t contig_id;
  int32_t n_paths;       /* haplotypes AssemblyUtils.discoverHaplotypes returns: 0 = no path, or more than 16 */
  int64_t locus;
  int32_t n_graph_nodes; /* k-mers with count >= min_occurrence */
  int32_t consumed;      /* 1: the reference's walk visits this window; 0: assembled speculatively, then skipped by
                            the window jump of an earlier successful assembly */
} guac_assembled_window;

/*
 * germline-assembly: windowFlatMapWithState(halfWindowSize = assembly_window_range, skipEmpty = true) over one
 * sample with the closure of GermlineAssemblyCaller.discoverGermlineVariants
 * (M/commands/GermlineAssemblyCaller.scala:84-219): loci with fewer than min_alt_reads non-reference reads are
 * skipped; with shortcut_assembly, loci whose reads are not an active region (M/assembly/AssemblyUtils.scala:28-39)
 * go to the toy caller (guac_germline_standard); the others are assembled — De Bruijn graph of the window's reads
 * (M/assembly/DeBruijnGraph.scala:342-373), unipath merge (:91-126), depth-first search from the first to the last
 * reference k-mer (:183-300), paths scored against the reads when there are more than two
 * (M/assembly/AssemblyUtils.scala:50-114), each kept path aligned to the reference slice
 * (M/alignment/ReadAlignment.scala:64-195) and turned into variants (AssemblyUtils.scala:127-174).
 * Assembled calls carry n_genotypes = 0 and the window-wide evidence of :162-192.  The partition state
 * (lastCalledLocus, the jump by range - kmer_size) is that of one partition per contig: call once per
 * LociPartitioning partition to reproduce a partitioned run.
 */
int guac_germline_assembly(guac_ctx* ctx, const guac_ga_params* params, const guac_range* loci, int32_t n_ranges,
                           guac_called_allele** out_calls, int64_t* n_calls, char** out_arena, int64_t* arena_len,
                           guac_assembled_window** out_windows, int64_t* n_windows, int64_t* out_n_visited);

/*
 * DeBruijnGraph.discoverPathsFromReads (M/assembly/DeBruijnGraph.scala:437-464) over every uploaded read of one
 * sample — graph with mergeNodes = true, depthFirstSearch from the first to the last kmer_size bases of
 * reference_sequence — followed by mergeOverlappingSequences of each path (M/assembly/AssemblyUtils.scala:84).
 * out_paths receives the haplotypes in the reference's list order (most recently found first), each NUL-terminated;
 * max_paths <= 17.  This is the assembler of guac_germline_assembly applied to a caller-chosen read set.
 */
int guac_discover_paths(guac_ctx* ctx, int32_t sample_id, const char* reference_sequence, int32_t reference_len, int32_t kmer_size,
                        int32_t min_occurrence, int32_t max_paths, int32_t min_mean_kmer_quality, char** out_paths, int64_t* out_len,
                        int32_t* n_paths);

/*
 * ReadAlignment(sequence, reference, mismatchProbability, openGapProbability, closeGapProbability)
 * (M/alignment/ReadAlignment.scala:64-90, scoreAlignmentPaths :92-195): the best affine-gap alignment of the whole
 * sequence to any stretch of the reference; a probability <= 0 selects the reference's default (e^-4, e^-6, 1 - e^-1).
 * out_cigar receives toCigarString (:22-43, operators = X I D).
 */
int guac_read_alignment(guac_ctx* ctx, const char* sequence, int32_t sequence_len, const char* reference, int32_t reference_len,
                        double mismatch_probability, double open_gap_probability, double close_gap_probability, char* out_cigar,
                        int32_t cigar_cap, int32_t* ref_start, int32_t* ref_end, int32_t* alignment_score, double* raw_score);

/* ---- somatic-joint ----------------------------------------------------------
 * M/jointcaller/Parameters.scala:10-25 plus the inputs (M/jointcaller/Input.scala, InputCollection.scala:55-101)
 * and the flags of M/commands/SomaticJointCaller.scala:154-163. */
enum { GUAC_SJ_NORMAL_DNA = 0, GUAC_SJ_TUMOR_DNA = 1, GUAC_SJ_TUMOR_RNA = 2 };
#define GUAC_SJ_MAX_SAMPLES 16
#define GUAC_SJ_MAX_DEPTH_ENTRIES 8
typedef struct {
  int32_t n_samples;                          /* = n_samples of the uploaded batch */
  int32_t sample_kind[GUAC_SJ_MAX_SAMPLES];   /* GUAC_SJ_* per sample id (tissue type x analyte) */
  int32_t max_alleles_per_site;               /* 4 (<= GUAC_SJ_MAX_DEPTH_ENTRIES - 1) */
  int32_t any_allele_min_supporting_reads;    /* 2 */
  double any_allele_min_supporting_percent;   /* 2.0 */
  double germline_neg_log10_het_prior;        /* 4 */
  double germline_neg_log10_hom_alt_prior;    /* 5 */
  double somatic_neg_log10_variant_prior;     /* 6 */
  double somatic_neg_log10_variant_prior_rna; /* 1 */
  double somatic_neg_log10_variant_prior_with_rna_evidence; /* 1 */
  double somatic_vaf_floor;                   /* 0.05 */
  double filter_strand_bias_phred;            /* 60 */
  double filter_somatic_normal_nonreference_percent; /* 3.0 */
  int32_t filter_somatic_normal_depth;        /* 30 */
  int32_t only_somatic;
  int32_t include_filtered;
  int32_t _pad;
} guac_sj_params;

typedef struct {
  uint32_t seq_off, seq_len; /* allele string in the arena */
  int32_t count;
  int32_t _pad;
} guac_sj_depth_entry;

/* One SingleSampleSingleAlleleEvidence (M/jointcaller/evidence/{NormalDNA,TumorDNA,TumorRNA}SingleSampleSingleAlleleEvidence.scala)
 * with its annotations (M/jointcaller/annotation/StrandBias.scala:22-58). */
typedef struct {
  int32_t kind;              /* GUAC_SJ_* */
  int32_t total_depth;       /* PileupStats.totalDepthIncludingReadsContributingNoAlleles */
  int32_t n_depths;          /* truncatedAllelicDepths(maxAllelesPerSite + 1), most frequent first */
  int32_t n_likelihoods;     /* 3 for normal DNA: ref/ref, ref/alt, alt/alt; 2 for tumor: {ref}, {alt: vaf, ref: 1 - vaf} */
  guac_sj_depth_entry depths[GUAC_SJ_MAX_DEPTH_ENTRIES];
  double log_likelihoods[3]; /* log10 */
  double posteriors[3];      /* likelihood - prior */
  double alt_vaf;            /* tumor: max(somaticVafFloor, vaf(alt)) */
  int32_t sb_variant_forward, sb_variant_reverse, sb_total_forward, sb_total_reverse;
  double sb_phred;
  int32_t sb_filtered;
  int32_t triggered;         /* tumor DNA: sample triggered the call; tumor RNA: expressed */
} guac_sj_sample_evidence;

/* One MultiSampleSingleAlleleEvidence (M/jointcaller/evidence/MultiSampleSingleAlleleEvidence.scala:24-342).
 * Its per-sample evidences are sample_evidence[first_sample_evidence .. + n_samples + 2):
 * the samples in order, then pooled normal DNA, then pooled tumor DNA. */
typedef struct {
  int32_t contig_id;
  int32_t site_index;        /* evidences of one MultiSampleMultiAlleleEvidence share it */
  int64_t start;             /* allele.start (0-based) = pileup locus + 1 */
  uint32_t ref_off, ref_len, alt_off, alt_len;
  int32_t is_germline_call, is_somatic_call;
  int32_t germline_genotype; /* argmax of the pooled germline posteriors: 0 ref/ref, 1 ref/alt, 2 alt/alt */
  int32_t tumor_rna_expressed;
  int32_t has_insufficient_normal, in_reference_reads, in_total_reads, in_filtered;
  uint32_t failing_filters;  /* bit 0 STRAND_BIAS, bit 1 INSUFFICIENT_NORMAL */
  int32_t forced;            /* site was force-called */
  int64_t first_sample_evidence;
} guac_sj_allele_evidence;

/*
 * somatic-joint: SomaticJoint.makeCalls (M/commands/SomaticJointCaller.scala:154-203) =
 * pileupFlatMapMultipleSamples(skipEmpty = true, MultiSampleMultiAlleleEvidence.make) over the loci shifted
 * by -1 (lociSetMinusOne :143-152; the function does the shift itself).  A site is force-called when
 * pileup locus + 1 lies in `force_loci` (:181).  Per-element work — ReadSubsequence, AlleleAtLocus,
 * PileupStats and the mixture likelihoods (M/jointcaller/pileup_summarization/ (all files), AlleleAtLocus.scala:85-149,
 * evidence/ ...SingleSampleSingleAlleleEvidence.scala) — and posteriors, call logic, annotations
 * (M/jointcaller/evidence/MultiSampleSingleAlleleEvidence.scala:66-341, annotation/StrandBias.scala:47-58 over
 * filters/FishersExactTest.scala:12-19, annotation/InsufficientNormal.scala:21-53) and the emit policy
 * (MultiSampleMultiAlleleEvidence.scala:131-177) all run on the device: only the emitted records come back.
 * Output: one guac_sj_allele_evidence per emitted allele, in (range order, locus, allele) order, each with
 * n_samples + 2 guac_sj_sample_evidence records; strings in the arena.
 */
int guac_somatic_joint(guac_ctx* ctx, const guac_sj_params* params, const guac_range* loci, int32_t n_ranges,
                       const guac_range* force_loci, int32_t n_force, guac_sj_allele_evidence** out_alleles, int64_t* n_alleles,
                       guac_sj_sample_evidence** out_samples, int64_t* n_sample_evidences, char** out_arena, int64_t* arena_len,
                       int64_t* out_n_visited);

/* ---- other consumers of the pileup operator ------------------------------------
 * M/commands/VariantSupport.scala:100-112 (pileupsToAlleleCounts through pileupFlatMapMultipleSamples, skipEmpty = true):
 * one record per (visited locus, sample, distinct allele of the pileup's elements) with the number of elements that
 * carry it.  Alleles as M/pileup/Alignment.scala:14-75 defines them — Match (b, b), Mismatch (ref, b), Insertion
 * (sequenced head, head + inserted), Deletion (ref incl. deleted, ref head), MidDeletion (ref, ""), Clipped ("", "").
 * Records come in (range order, locus), sample, (ref, alt) string order (the reference's order inside a pileup is a
 * hash map's; its suite sorts before comparing). */
typedef struct {
  int32_t contig_id;
  int32_t sample_id;
  int64_t locus;
  uint32_t ref_off, ref_len, alt_off, alt_len; /* into the returned string arena */
  int32_t count;
  int32_t _pad;
} guac_allele_count;
int guac_variant_support(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, guac_allele_count** out_counts, int64_t* n_out,
                         char** out_arena, int64_t* arena_len);

/* M/commands/VAFHistogram.scala:30-58 (VariantLocus) */
typedef struct {
  int32_t sample_id;
  int32_t contig_id;
  int64_t locus;
  float variant_allele_frequency; /* (depth - referenceDepth).toFloat / depth */
  int32_t _pad;
} guac_variant_locus;
/*
 * VAFHistogram.variantLociFromReads + generateVAFHistograms (M/commands/VAFHistogram.scala:194-245): every (locus, sample)
 * with depth >= min_read_depth, referenceDepth != depth and variant allele frequency >= min_vaf_percent / 100.0 is a
 * VariantLocus; out_hist[s * 101 + b] counts sample s's loci whose rounded percentage falls in bin b = p - p % (100 / bins)
 * (p = (vaf * 100).toInt), out_n_variant_loci[s] their number.  Both are caller-provided ([n_samples * 101], [n_samples]).
 * out_loci (optional, may be NULL) receives the records in (range order, locus, sample) order.
 */
int guac_vaf_histogram(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, int32_t bins, int32_t min_read_depth,
                       int32_t min_vaf_percent, int64_t* out_hist, int64_t* out_n_variant_loci, guac_variant_locus** out_loci,
                       int64_t* n_loci);

/* ---- coverage-aware loci partitioning -------------------------------------------
 * The "capped" LociPartitioner (M/loci/partitioning/CappedRegionsPartitioner.scala:66-88, the reference's default):
 * CoverageRDD.traversalCoverage (M/readsets/rdd/CoverageRDD.scala:120-143, M/readsets/iterator/ContigCoverageIterator.scala:28-117)
 * gives every eligible locus a Coverage(depth, starts) over the uploaded reads of all samples, a read counting at every
 * locus of [start - half_window, end + half_window); TakeLociIterator (M/loci/set/TakeLociIterator.scala:25-196) then cuts
 * greedy loci sets of at most max_reads_per_partition reads, dropping loci deeper than that.  `loci` must be in contig
 * order (the reference visits contigs in name order; here the caller's contig ids are the order).  Output: intervals
 * out_ranges[i] of loci set out_partition[i], in order; with trim_ranges = 0 one interval per (loci set, stretch of a
 * contig) spanning uncovered loci, as the reference's default.  The result feeds the shards of a multi-GPU run in place of
 * UniformPartitioner's ranges when coverage is uneven. */
int guac_capped_partitioning(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, int32_t half_window, int32_t max_reads_per_partition,
                             int32_t trim_ranges, guac_range** out_ranges, int32_t** out_partition, int64_t* n_out, int32_t* n_partitions);
/* TakeLociIterator alone, over (position, coverage) pairs the caller already has (M/loci/set/TakeLociIterator.scala:25-196;
 * the form TakeLociIteratorSuite feeds): entries sorted by (contig, locus), uncovered loci absent. */
int guac_take_loci(guac_ctx* ctx, const int32_t* contig_id, const int64_t* locus, const int32_t* depth, const int32_t* starts, int64_t n,
                   int32_t max_regions_per_partition, int32_t trim_ranges, guac_range** out_ranges, int32_t** out_partition, int64_t* n_out,
                   int32_t* n_partitions);

/* ---- measurement support ----------------------------------------------------
 * These have no counterpart in the reference: they exist so that bench.py and the
 * tests can generate SURVEY.md 8(d)'s seeded synthetic reads directly in HBM, read them
 * back, and time the kernels on the library's own stream. */

/* CUDA stream (cudaStream_t) every kernel of this context is launched on. */
void* guac_stream(guac_ctx* ctx);
/* Implementation switches for tests and A/B measurements:
 *   "ss_impl"               somatic-standard: 2 = read pass + locus pass as two kernels (default), 0 = both in one tile
 *                           kernel, 1 = generic gather kernel only; all three fall back to the gather kernel per tile
 *   "ss_force_fallback_mod" k > 0 sends every k-th tile through the fallback path (test hook)
 *   "sj_impl"               somatic-joint: 2 = read pass + classifying locus pass + fast site kernel (default), 1 = the same
 *                           with the single-kernel locus pass, 0 = every candidate locus through the generic per-locus
 *                           kernels; more than four samples always take the generic kernels
 *   "sj_host_threads"       accepted and ignored (the per-site arithmetic runs on the device)
 *   "sj_prune"              1 (default): the joint caller's locus pass drops candidate sites that certainly hold no call;
 *                           0: every candidate reaches the site kernels (test hook)
 *   "synth_profile"         guac_synth_generate: 0 (default) the workload of SURVEY.md 8(d); 1 its "realistic" variant (read
 *                           density varying like a Poisson pileup's, 15 % of reads with a quality-2 tail, 10 % multi-operator
 *                           reads, 1 % of reads with 10..20 mismatches)
 *   "compact_min_qual_bits" guac_reads_export_compact: 2 (default) the narrowest quality form the batch allows; 4 or 8 force
 *                           the wider ones (measurement hook: what real, unbinned qualities cost on the wire)
 *   "compact_narrow"        1 (default): guac_reads_export_compact emits the narrow columns where the batch allows; 0: never
 *   "compact_ref_relative"  1: guac_reads_export_compact emits reference-relative bases (mm_count / mm_words) when every contig
 *                           of the batch has its reference window uploaded; 0 (default): every read's bases in bases2
 *   "ss_rank_max"           somatic-standard ranks more calls than this through a bitmap over the loci instead of by
 *                           pairwise comparison (test hook; default and maximum 32768, negative restores the default) */
int guac_set_option(guac_ctx* ctx, const char* name, int64_t value);
/* Tiles of the most recent guac_somatic_standard call that took the fallback path. */
int64_t guac_last_fallback_tiles(guac_ctx* ctx);
/* Loci of that call whose likelihoods had to be evaluated in full (not excluded by the LOD bound). */
int64_t guac_last_heavy_loci(guac_ctx* ctx);
/* Counters of the most recent caller invocation by name: "fallback_tiles", "heavy_loci" (as above; for somatic-joint the
 * candidate loci that went to the generic per-locus kernels), "sj_simple" (candidate loci the fast site kernel took),
 * "sj_bail" (those it handed back), "sj_sites" (sites with at least one allele), "sj_emitted_sites".  -1: unknown name. */
int64_t guac_last_stat(guac_ctx* ctx, const char* name);
/* Device time in ms (CUDA events on the context's stream) of the dominant per-locus kernel
 * of the most recent caller invocation, and how many loci it processed. */
int guac_last_kernel_ms(guac_ctx* ctx, float* ms, int64_t* n_loci);

/* Generates the synthetic batch (guacamole_b200/synth.py is the specification) for one
 * contig directly on the device, as if guac_ref_upload + guac_reads_upload had been
 * called: reference of contig_len loci, sum(depths)*contig_len/100 reads of 100 bases,
 * samples interleaved proportionally to depths[0..n_samples), tumor[s] != 0 marks samples
 * that carry the planted somatic SNVs. */
int guac_synth_generate(guac_ctx* ctx, int32_t contig_id, int64_t contig_len, int32_t n_samples, const int32_t* depths,
                        const uint8_t* tumor, uint64_t seed);

/* Copies the uploaded (or generated) batch back into library-owned pinned host memory;
 * every pointer of *out is released with guac_free. */
int guac_reads_export(guac_ctx* ctx, guac_read_batch* out);
/*
 * Compact wire form of a read batch for hosts that ship a partition over PCIe every call: the same reads as
 * guac_read_batch with everything derivable left out — offsets become lengths (prefix sums run on the device),
 * the per-read contig id becomes runs, the not-ACGT mask is rebuilt from the exception list, and base qualities
 * travel as 2- or 4-bit codes into a table when the batch holds at most 4 / 16 distinct qualities (binned
 * Illumina qualities; otherwise one byte per base as before).  100-base single-operator reads: 58 / 83 B per read with
 * the narrow columns at the end of the struct (69 / 94 B without) instead of 168.5.  Lossless; replaces the same `Iterator[MappedRead]` as guac_reads_upload.
 */
typedef struct {
  int64_t n_reads;
  int32_t n_samples;
  int32_t n_contig_runs;
  const int32_t* run_contig_id;   /* [n_contig_runs] contig of each run of consecutive reads */
  const int64_t* run_first_read;  /* [n_contig_runs + 1] first read of each run; last entry = n_reads */
  const int64_t* start;           /* [n_reads] */
  const uint16_t* read_len;       /* [n_reads] sequenced bases (<= 65535) */
  const uint16_t* n_cigar_ops;    /* [n_reads] */
  const uint32_t* cigar;          /* BAM-encoded operators of all reads, concatenated */
  const uint8_t* bases2;          /* 2 bit/base, concatenated, as in guac_read_batch */
  int64_t n_exc;
  const uint64_t* exc_pos;        /* sorted batch-global base indices of the bases that are not A C G T */
  const uint8_t* exc_base;
  int32_t qual_bits;              /* 8: one byte per base; 4 / 2: two / four bases per byte, lowest bits = lowest base index */
  int32_t _pad;
  uint8_t qual_table[16];         /* qual_bits < 8: code -> quality */
  const uint8_t* qual;
  const uint8_t* mapq;            /* [n_reads] */
  const uint8_t* flags;           /* [n_reads] */
  const uint8_t* sample_id;       /* [n_reads] */
  /* Narrower forms of three columns (each optional: NULL / 0 = the wide column above is given instead).
   * 100-base single-operator reads then travel in 58 B per read. */
  const int64_t* run_start_base;  /* [n_contig_runs], with start_rel */
  const uint32_t* start_rel;      /* [n_reads] start = run_start_base[run of the read] + start_rel; replaces `start` */
  const uint8_t* n_cigar_ops8;    /* [n_reads] 0: the CIGAR is one M operator over all the read's bases and has no word in
                                   * `cigar`; k > 0: the read's k words in `cigar`.  Replaces `n_cigar_ops` */
  int32_t uniform_read_len;       /* > 0: every read has this many bases; replaces `read_len` */
  int32_t _pad2;
  /* Reference-relative bases (optional; NULL = `bases2` holds every read).  A read with one aligned operator (M, = or X over
   * all its bases), only A C G T bases, lying wholly on A C G T reference inside the uploaded window of its contig, and
   * shorter than 16384 bases, can travel as its differences from the reference: mm_count[read] = their number (< 255) and one
   * word each in `mm_words`, read after read: position in the read << 2 | (reference code XOR read code), codes A C G T =
   * 0 1 2 3.  mm_count = 255: the read's bases are in `bases2`, which then holds those reads only, concatenated 2 bits per
   * base as before (n_explicit_bases of them).  The reference windows of the batch's contigs must be uploaded before the
   * reads.  100-base reads of a 0.4 % error rate: 58 -> about 35 B per read (the pileup walks every base of
   * `MappedRead.sequence`, M/reads/MappedRead.scala:28-35, whichever way it arrived). */
  const uint8_t* mm_count;        /* [n_reads] */
  const uint16_t* mm_words;       /* [n_mm_words] */
  int64_t n_mm_words;
  int64_t n_explicit_bases;
} guac_read_batch_compact;

int guac_reads_upload_compact(guac_ctx* ctx, const guac_read_batch_compact* batch);
/* The uploaded (or generated) batch in the compact form, in library-owned pinned host memory (every pointer is
 * released with guac_free).  Measurement / test support. */
int guac_reads_export_compact(guac_ctx* ctx, guac_read_batch_compact* out);

/* Copies [offset, offset+len) of a contig's reference window to `dst` (host). */
int guac_ref_export(guac_ctx* ctx, int32_t contig_id, int64_t offset, int64_t len, uint8_t* dst);

/* ---- host stages either side of the device path (no GPU work; usable without a context) -------------------
 * One partition's decoded SAM records, column-wise, as a JNI shim copies them out of htsjdk SAMRecords — the inputs of
 * Read.apply (M/reads/Read.scala:63-116). */
typedef struct {
  int64_t n_records;
  int32_t n_samples;
  int32_t _pad;
  const uint16_t* sam_flags;       /* [n] SAM FLAG: 0x1 paired, 0x4 unmapped, 0x10 reverse strand, 0x200 failed QC, 0x400 duplicate */
  const int32_t* reference_index;  /* [n] getReferenceIndex; < 0: none (the contig ids of this library) */
  const int64_t* alignment_start;  /* [n] getAlignmentStart, 1-based */
  const uint8_t* mapping_quality;  /* [n] */
  const uint32_t* cigar_off;       /* [n + 1] */
  const uint32_t* cigar;           /* BAM-encoded operators */
  const uint64_t* seq_off;         /* [n + 1] */
  const uint8_t* bases_ascii;      /* getReadString bytes */
  const uint8_t* base_qualities;   /* getBaseQualities (Phred, not +33) */
  const uint8_t* sample_id;        /* [n] */
} guac_sam_records;

/* InputConfig (M/readsets/io/InputConfig.scala:20-45) as ReadSets.filterRDD applies it (M/readsets/ReadSets.scala:272-304) */
typedef struct {
  int32_t non_duplicate;                /* drop reads with the duplicate bit */
  int32_t passed_vendor_quality_checks; /* drop reads with the failed-QC bit */
  int32_t is_paired;                    /* keep paired-end reads only */
  int32_t min_alignment_quality;        /* > 0: keep mapped reads with at least this mapping quality */
  const guac_range* overlaps_loci;      /* n_overlaps_loci > 0: keep reads that overlap one of these ranges */
  int32_t n_overlaps_loci;
  int32_t _pad;
} guac_read_filters;

/*
 * Read.apply + ReadSets.filterRDD + the partition's (contig, start) order (M/readsets/rdd/PartitionedRegions.scala:186-208)
 * + the packer: mapped records that pass the filters become one guac_read_batch_compact (narrow columns and 2- / 4-bit
 * quality codes where the reads allow them), built on n_threads host threads (<= 0: one per core).  With a context the
 * columns live in its pinned pool (release each with guac_free); with ctx == NULL on the C heap (guac_host_free).
 * Feed the result to guac_reads_upload_compact.
 */
int guac_pack_reads(guac_ctx* ctx, const guac_sam_records* records, const guac_read_filters* filters, int32_t n_threads,
                    guac_read_batch_compact* out, int64_t* n_kept);
/* One contig window of the reference in host memory (ASCII upper-case, 0 = missing), as ReferenceBroadcast holds it
 * (M/reference/ReferenceBroadcast.scala:70-83). */
typedef struct {
  int32_t contig_id;
  int32_t _pad;
  int64_t offset;        /* locus of bases[0] */
  int64_t len;
  const uint8_t* bases;
} guac_host_reference;
/* guac_pack_reads with the reference-relative base column (guac_read_batch_compact::mm_count): a read of one aligned operator
 * over A C G T bases that lies wholly on A C G T bases of its contig's window travels as its differences from it.  The same
 * windows must be on the device (guac_ref_upload) before the result is uploaded. */
int guac_pack_reads_ref(guac_ctx* ctx, const guac_sam_records* records, const guac_read_filters* filters, const guac_host_reference* refs,
                        int32_t n_refs, int32_t n_threads, guac_read_batch_compact* out, int64_t* n_kept);
void guac_host_free(void* p);

/* SomaticGenotypeFilter.SomaticGenotypeFilterArguments (M/filters/somatic/SomaticGenotypeFilter.scala:34-80); < 0 = not set */
typedef struct {
  int32_t min_likelihood;                 /* --min-likelihood, not set */
  int32_t min_vaf;                        /* --min-vaf, 0 */
  int32_t min_average_mapping_quality;    /* 20 */
  int32_t min_average_base_quality;       /* 20 */
  int32_t min_tumor_read_depth;           /* 5 */
  int32_t min_normal_read_depth;          /* 5 */
  int32_t max_tumor_read_depth;           /* Int.MaxValue */
  int32_t min_tumor_alternate_read_depth; /* not set */
  int32_t maximum_median_mismatches;      /* not set */
  int32_t _pad;
} guac_ss_filter_params;
/* SomaticGenotypeFilter.apply (M/filters/somatic/SomaticGenotypeFilter.scala:82-130, with the reference's quirk that the
 * "average base quality" filter tests meanMappingQuality): keep[i] = 1 for the calls that pass. */
int guac_ss_genotype_filter(const guac_ss_filter_params* params, const guac_ss_call* calls, int64_t n, uint8_t* keep);
/* The header line and data rows the reference's VCF writer renders for the (kept) calls
 * (M/variants/CalledSomaticAllele.scala:45-64, M/variants/GenotypeOutputArgs.scala:86-113), '\n' separated.  Text functions
 * write into the caller's buffer: *needed always receives the length; GUAC_E_OOM when cap is too small (call again). */
int guac_ss_vcf_rows(const guac_ss_call* calls, int64_t n, const uint8_t* keep /* may be NULL: all */, const char* arena,
                     const char* const* contig_names, const char* sample_name, char* out_text, int64_t cap, int64_t* needed);
/* VCFOutput.writeVcf's rows for guac_somatic_joint's records (M/jointcaller/VCFOutput.scala:97-100,127-413 through htsjdk
 * 2.6.1): include_pooled_* < 0 = the caller's default (more than one sample of the kind, SomaticJointCaller.scala:223);
 * somatic_genotype_policy 0 = presence (default), 1 = trigger. */
int guac_sj_vcf_rows(const guac_sj_params* params, const guac_sj_allele_evidence* alleles, int64_t n_alleles,
                     const guac_sj_sample_evidence* samples, const char* arena, const char* const* contig_names,
                     const char* const* sample_names, int32_t include_pooled_normal, int32_t include_pooled_tumor,
                     int32_t somatic_genotype_policy, char* out_text, int64_t cap, int64_t* needed);
/* Number formatting as the JVM does it (HALF_UP on the exact decimal expansion): kind 0 String.format("%.8g"), 1 htsjdk
 * VCFEncoder.formatVCFDouble, 2 String.format("%.2f").  Test support for the two row writers. */
int guac_format_double(int32_t kind, double x, char* out, int32_t cap);

#ifdef __cplusplus
}
#endif
#endif /* GUAC_B200_H */
