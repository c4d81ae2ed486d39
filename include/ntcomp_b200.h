/* ntcomp_b200.h -- C ABI of the B200-native ntcomp encode/decode hot path (libntcomp_b200.so).
 *
 * The boundary a host of tmaklin/ntcomp's sequence codec binds (reference file:line in each comment).  Results are
 * bit-exact with this repository's CPU oracle (oracle/), a statement-by-statement restatement of src/lib.rs, encode.rs and
 * decode.rs over restated crate semantics; equality with the Rust binary itself is NOT yet pinned (no toolchain where this
 * was built; DESIGN.md section 2, INTEGRATION.md section 0 and rust/ref_dump say how to pin it).
 * Plain pointers and sizes only; every call returns an int32 status (NTC_OK == 0) and never
 * throws/panics across the boundary; ntc_last_error() returns the calling thread's last message.
 * One ntc_ctx per GPU, used from one host thread at a time; different contexts are independent.
 * There is NO CPU fallback: every compute entry point fails with NTC_ERR_CUDA without a GPU.
 *
 * Record word (src/encode.rs:145-158, little-endian u64):
 *   reference record (len > 11): bytes 0-3 colex rank (u32), bytes 4-6 len (u24), byte 7 = first (0/1)
 *   literal record   (len <= 11): bytes 0-6 2-bit bases (base j at bits 2j..2j+1, A0 C1 G2 T3),
 *                                 byte 7 = (first + 2) | len << 2
 * Records of a read run from the read's END to its START; `first` marks the END segment.
 */
#ifndef NTCOMP_B200_H
#define NTCOMP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ntc_index ntc_index; /* host-side flat SBWT + LCS (replaces SbwtIndexVariant + LcsArray, src/lib.rs:20-22) */
typedef struct ntc_ctx ntc_ctx;     /* one GPU: re-laid-out index in HBM + workspaces */

enum {
    NTC_OK = 0,
    NTC_ERR_INVALID_ARG = 1,
    NTC_ERR_CUDA = 2,        /* no device / CUDA failure: there is no CPU fallback */
    NTC_ERR_CAPACITY = 3,    /* caller's output buffer too small */
    NTC_ERR_EMPTY_READ = 4,  /* EncodeError, src/lib.rs:224 / src/encode.rs:133 */
    NTC_ERR_BAD_BASE = 5,    /* non-ACGT byte, or a base absent from the index (d == 0): the reference
                                underflows at src/lib.rs:207 */
    NTC_ERR_LEN24 = 6,       /* a record length >= 2^24, src/lib.rs:226 */
    NTC_ERR_IO = 7,
    NTC_ERR_BAD_RECORD = 8,  /* colex rank >= n_sets in a record handed to decode */
    NTC_ERR_UNSUPPORTED = 9, /* k outside [12,255] or n_sets >= 2^32 (src/encode.rs:147 truncates to u32, Q2) */
    NTC_ERR_EMPTY_COLUMN = 10 /* a block without a literal or without a reference record cannot be coded: the
                                reference's write_block_to returns Err and main.rs:170 drops the block (Q4) */
};

const char* ntc_last_error(void);
const char* ntc_version(void);

/* ---------------- index (host) ------------------------------------------------------------- */

/* Host SBWT+LCS construction; replaces kbo::build(&seqs, opts) as called at src/main.rs:111-134 and
 * tests/fasta_data.rs:49-56 (add_revcomp = true, build_select = true there).  Offline preprocessing. */
int32_t ntc_index_build(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, uint32_t k,
                        int32_t add_revcomp, int32_t n_threads, ntc_index** out);

/* The same construction on the GPU ("next" row f4: `ntcomp build`, src/main.rs:97-140): both strands of all sequences become
 * one 2-bit text in HBM, the k-mer occurrences are radix-sorted by their k-mer in colex order (4 characters per pass, gathered
 * from the text -- keys are never materialised), distinct k-mers, sources, '$'-padded dummy nodes, LCS and the subset rows
 * follow from scans, binary searches and one merge.  Same index, bit for bit, as ntc_index_build.  No CPU fallback.
 * stats (nullable, 8 doubles): [0] ms until the input is in HBM, [1] ms of device work + download, [2] k-mer occurrences,
 * [3] distinct k-mers, [4] source k-mers, [5] dummy nodes, [6] radix passes, [7] kernel launches. */
int32_t ntc_index_build_gpu(int32_t device, const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, uint32_t k,
                            int32_t add_revcomp, ntc_index** out, double* stats);

/* Flattened view of an existing SbwtIndex<SubsetMatrix> + LcsArray (what a Rust host passes after
 * kbo::index::load_sbwt, src/main.rs:149,190): 4 subset rows of n_sets bits (LSB-first u64 words,
 * A,C,G,T) and LCS as one byte per node.  C is recomputed from the rows. */
int32_t ntc_index_from_arrays(uint32_t k, uint64_t n_sets, const uint64_t* row_a, const uint64_t* row_c,
                              const uint64_t* row_g, const uint64_t* row_t, const uint8_t* lcs, ntc_index** out);

/* Own flat file format (stands in for kbo::index::serialize_sbwt / load_sbwt, src/main.rs:138,149). */
int32_t ntc_index_save(const ntc_index* idx, const char* path);
int32_t ntc_index_load(const char* path, ntc_index** out);

uint32_t ntc_index_k(const ntc_index* idx);          /* sbwt.k(), src/lib.rs:171 */
uint64_t ntc_index_n_sets(const ntc_index* idx);
void ntc_index_get_C(const ntc_index* idx, uint64_t out[5]);
void ntc_index_get_row(const ntc_index* idx, int32_t c, uint64_t* out /* ceil(n_sets/64) words */);
void ntc_index_get_lcs(const ntc_index* idx, uint8_t* out /* n_sets bytes */);
void ntc_index_free(ntc_index* idx);

/* ---------------- context (one per GPU) ---------------------------------------------------- */

int32_t ntc_ctx_create(int32_t device, const ntc_index* idx, ntc_ctx** out);
/* The same, holding only the derived arrays one direction reads (the CLI's encode / decode subcommands; the
 * reference loads one index for either, src/main.rs:138,185): NTC_USE_ENCODE = rank blocks + LCS words + prefix
 * table, NTC_USE_DECODE = jump table.  Entry points of the other direction fail with NTC_ERR_UNSUPPORTED. */
#define NTC_USE_ENCODE 1u
#define NTC_USE_DECODE 2u
int32_t ntc_ctx_create_for(int32_t device, const ntc_index* idx, uint32_t uses, ntc_ctx** out);
void ntc_ctx_destroy(ntc_ctx* ctx);
/* bytes of HBM held by the re-laid-out index: [0]=rank blocks, [1]=LCS, [2]=decode jump table */
void ntc_ctx_index_bytes(const ntc_ctx* ctx, uint64_t out[3]);

/* Copy of one derived device array as the kernels see it (which: 0 rank blocks, 1 LCS words, 2 decode jump table, 3 prefix table;
 * *n_bytes = its size, 0 when this context does not hold it; out may be NULL to ask for the size).  Diagnostic: the tests compare the
 * arrays built by kernels at context creation with the host builders' (NTC_HOST_LAYOUTS=1 makes ntc_ctx_create use those). */
int32_t ntc_ctx_layout_read(ntc_ctx* ctx, int32_t which, void* out, uint64_t cap, uint64_t* n_bytes);

/* Matching-statistics tile length.  A read longer than the tile is processed as independent tiles of
 * that many positions, each warmed up on the k-1 bases before it (long reads, BASELINE config C5: a
 * 10 kbp read is ~20 chains instead of one); the output is identical for every tile length.
 * 0 (default) = chosen per batch; otherwise >= 256, rounded up to a multiple of 4. */
int32_t ntc_ctx_set_tile_len(ntc_ctx* ctx, uint32_t tile_len);

/* Host-buffer entry points: serial == 0 (default) moves results back while the next chunk of input goes up (both PCIe
 * directions busy); serial != 0 sends all input first and the results afterwards.  Same results either way.  Meant for hosts
 * where mixed-direction DMA is slower than either direction alone (8 GPUs copying at once on the round-2 box: 29.8 GB/s up
 * alone, 14.9 down alone, 19.3 for both together); measured there it gains nothing because the ranks are out of phase
 * (profiles/r2/ab_copy_directions.txt), so the default is duplex.  NTC_COPY_DIRECTIONS=serial|duplex sets the initial value. */
int32_t ntc_ctx_set_copy_directions(ntc_ctx* ctx, int32_t serial);

/* Host threads of the host-buffer entry points.  PAGEABLE caller buffers (a Rust Vec, a numpy array) are moved through page-locked
 * staging owned by the context, filled and emptied by a pool of host threads: the reads are packed to 2 bits per base on the way up,
 * records, offsets and text are copied.  A pageable cudaMemcpy runs at 7.5 GB/s up and 14 GB/s down on the B200 boxes; with the staging,
 * 1.5 G bases encode in 41 instead of 167-200 ms and decode to FASTA in 46 instead of 109-119 ms (16 threads; page-locked buffers: 32 + 31 ms;
 * profiles/r2/pageable_probe*.json).  Page-locked caller buffers (ntc_pinned_alloc, cudaHostAlloc, cudaHostRegister) are copied by the DMA
 * engines directly.  n_threads < 0 (the default) = the above with this GPU's share of the host's threads (at most 16, off below 2);
 * 0 = never use host threads; n > 0 = n threads, and reads in page-locked buffers are packed on the host too (measured no faster there:
 * profiles/r2/ab_host_pack.txt).  Same results in every mode.  NTC_HOST_THREADS sets the initial value. */
int32_t ntc_ctx_set_host_threads(ntc_ctx* ctx, int32_t n_threads);

/* ---------------- encode ------------------------------------------------------------------- */

/* flat_map(encode_dictionary . encode_sequence) over a batch of reads: src/main.rs:166,169 with
 * src/lib.rs:163-230 and src/encode.rs:129-166.  `bases` = reads concatenated (ASCII ACGT),
 * `offsets` = n_reads+1 byte offsets.  Writes the u64 records of all reads in input order and, if
 * rec_offsets != NULL, n_reads+1 record offsets.  HOST buffers. */
int32_t ntc_encode_batch(ntc_ctx* ctx, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads,
                         uint64_t* records, uint64_t cap, uint64_t* n_records, uint64_t* rec_offsets);

/* encode_sequence alone (src/lib.rs:163): per record the tuple (len, interval.start, interval.end)
 * as three u64; `triples` holds 3*cap words. */
int32_t ntc_encode_sequence_batch(ntc_ctx* ctx, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads,
                                  uint64_t* triples, uint64_t cap, uint64_t* n_records, uint64_t* rec_offsets);

/* encode_dictionary alone (src/encode.rs:129): one read's dictionary of n (len,start,end) triples
 * -> n record words (literal bases are rebuilt from the index, as access_kmer does at :151). */
int32_t ntc_encode_dictionary(ntc_ctx* ctx, const uint64_t* triples, uint64_t n, uint64_t* records);

/* worst-case number of records for a batch with this many bases (one record per base) */
uint64_t ntc_encode_bound(uint64_t n_bases);

/* ---------------- decode ------------------------------------------------------------------- */

/* decode_sequence (src/lib.rs:254-318) over a record stream: sequences concatenated into `bases`,
 * n_seqs+1 offsets into seq_offsets (capacity cap_seqs+1 entries).  HOST buffers. */
int32_t ntc_decode_batch(ntc_ctx* ctx, const uint64_t* records, uint64_t n_records, uint8_t* bases, uint64_t cap,
                         uint64_t* seq_offsets, uint64_t cap_seqs, uint64_t* n_seqs);

/* decode_sequence + the FASTA writer loop of src/main.rs:201-209: ">seq.{N}\n{bases}\n" with
 * N = first_seq_number, first_seq_number+1, ...  (1-based, running across blocks). */
int32_t ntc_decode_fasta(ntc_ctx* ctx, const uint64_t* records, uint64_t n_records, uint64_t first_seq_number,
                         uint8_t* text, uint64_t cap, uint64_t* n_bytes, uint64_t* n_seqs);

/* exact output sizes for a record stream (host scan of the flag/len fields) */
void ntc_decode_bound(const uint64_t* records, uint64_t n_records, uint64_t first_seq_number, uint64_t* n_bases,
                      uint64_t* n_seqs, uint64_t* fasta_bytes);

/* ---------------- block container, host side ("next" row f1: src/lib.rs:232-252,320-363) ----------
 * ntc_block_write == write_block_to: splits the records into the four columns, codes them (minimal
 * binary, Rice, Rice, minimal binary), gzips each and prepends the 32-byte headers.  DEFLATE bytes
 * come from the system zlib (the reference links zlib-rs): members inflate to identical words.
 * ntc_block_read == decode_block up to (not including) decode_sequence; *consumed == 0 at EOF. */
uint64_t ntc_block_bound(uint64_t n_records);
int32_t ntc_block_write(const uint64_t* records, uint64_t n_records, uint64_t num_reads, uint8_t* out, uint64_t cap,
                        uint64_t* n_bytes);
int32_t ntc_block_read(const uint8_t* data, uint64_t size, uint64_t* records, uint64_t cap, uint64_t* n_records,
                       uint64_t* consumed);
/* ntc_block_read with options.  NTC_BLOCK_FULL_LAST_LITERAL_WORD: a block whose literal bases are a non-zero
 * multiple of 31 ends in a full literal word; the reference's zip_block_contents takes that word to hold
 * `bitnuc_len % 31` = 0 bases and panics at src/decode.rs:138, so ntc_block_read reports
 * NTC_ERR_IO for such a block.  With the flag the word is read as the 31 bases the encoder put there --
 * the file pipeline uses it, so every file the encoder writes can be decoded. */
#define NTC_BLOCK_FULL_LAST_LITERAL_WORD 1
int32_t ntc_block_read_ex(const uint8_t* data, uint64_t size, uint64_t* records, uint64_t cap, uint64_t* n_records,
                          uint64_t* consumed, int32_t flags);
/* write_block_to over HOST records with the column split and the Rice / minimal-binary bit packing on the GPU (values
 * scattered to prefix-summed ranks, one thread per output word gathers and assembles it: no per-value atomics); gzip on
 * the host.  Same bytes as ntc_block_write. */
int32_t ntc_block_write_gpu(ntc_ctx* ctx, const uint64_t* records, uint64_t n_records, uint64_t num_reads, uint8_t* out,
                            uint64_t cap, uint64_t* n_bytes);
/* decode_block up to decode_sequence (src/lib.rs:320-363) with the Rice / minimal-binary decoders (src/decode.rs:51-77)
 * and zip_block_contents (src/decode.rs:102-149) on the GPU: chunked speculative decoding of the code streams
 * (entry-state tables per chunk, composed by one warp per stream); gunzip on the host.  Same records, same
 * errors as ntc_block_read_ex. */
int32_t ntc_block_read_gpu(ntc_ctx* ctx, const uint8_t* data, uint64_t size, uint64_t* records, uint64_t cap,
                           uint64_t* n_records, uint64_t* consumed, int32_t flags);

/* ---- fused, device-resident column codec ("next" row f1 as SURVEY 8(f)1 defines it) ------------------------
 * The u64 records never leave HBM: the encoder's output goes straight into the column split and the integer coders
 * (src/encode.rs:59-94,168-229), and only the coded word streams + their parameters cross PCIe; the decoder takes
 * coded word streams and hands back FASTA text.  gzip / gunzip and the 32-byte headers stay on the host
 * (ntc_block_assemble / ntc_block_disassemble = the rest of write_block_to / decode_block). */
typedef struct ntc_coded_column {
    uint64_t n_values;     /* BlockHeader.num_u64, src/lib.rs:41 */
    uint64_t n_words;      /* BlockHeader.encoded_size: big-endian u64 words of the MSB-first code stream */
    uint64_t param;        /* BlockHeader.rice_param: Rice log2(b) or minimal-binary upper bound, src/encode.rs:117 */
    uint64_t word_offset;  /* first word of this column in the call's word array */
} ntc_coded_column;
typedef struct ntc_coded_block {
    uint64_t n_records;    /* records of the block (= n_values of the flags column) */
    uint64_t num_reads;    /* BlockHeader.num_records: the block-size argument of write_block_to (Q6) */
    int32_t status;        /* NTC_OK, or NTC_ERR_EMPTY_COLUMN: the reference drops this block (Q4) */
    int32_t reserved;
    ntc_coded_column col[4]; /* colex (minimal binary), len (Rice), flags (Rice), literal words (minimal binary): src/lib.rs:242-245 */
} ntc_coded_block;

/* flat_map(encode_dictionary . encode_sequence) over the reads + the column codec per block of block_reads reads
 * (65,536 at src/main.rs:152).  HOST reads in; per block the descriptor, and *words -> all coded words of the call in a
 * pinned buffer owned by ctx (valid until the next-but-one call on ctx). */
int32_t ntc_encode_columns(ntc_ctx* ctx, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, uint64_t block_reads,
                           ntc_coded_block* blocks, uint64_t cap_blocks, uint64_t* n_blocks, const uint64_t** words,
                           uint64_t* n_words, uint64_t* n_records);
/* four gzip members + 32-byte headers of one block (src/lib.rs:242-250, src/encode.rs:96-127); host only */
int32_t ntc_block_assemble(const ntc_coded_block* blk, const uint64_t* words, uint8_t* out, uint64_t cap, uint64_t* n_bytes);
/* One sub-block of the four (0 colex ranks, 1 match lengths, 2 flags, 3 literal words; src/lib.rs:242-250): header + gzip
 * member.  ntc_block_assemble's output is the concatenation of columns 0..3. */
int32_t ntc_sub_block_assemble(const ntc_coded_block* blk, int32_t column, const uint64_t* words, uint8_t* out, uint64_t cap, uint64_t* n_bytes);
/* four headers + gunzip of one block (src/lib.rs:326-355, src/decode.rs:79-91); host only.  *consumed == 0 at EOF. */
int32_t ntc_block_disassemble(const uint8_t* data, uint64_t size, ntc_coded_block* blk, uint64_t* words, uint64_t cap_words,
                              uint64_t* n_words, uint64_t* consumed);
/* decode_block (minus gunzip) + decode_sequence + the writer loop of src/main.rs:201-209 for a batch of blocks.  Every block
 * is decoded by itself as the reference does; numbering runs on from first_seq_number.  *text -> pinned buffer owned by
 * ctx (valid until the next-but-one call).  flags: NTC_BLOCK_FULL_LAST_LITERAL_WORD. */
int32_t ntc_decode_columns_fasta(ntc_ctx* ctx, const ntc_coded_block* blocks, uint64_t n_blocks, const uint64_t* words, uint64_t n_words,
                                 uint64_t first_seq_number, int32_t flags, const uint8_t** text, uint64_t* n_bytes, uint64_t* n_seqs,
                                 uint64_t* n_records);
/* one column through rice_encode / minimal_binary_encode (src/encode.rs:59-94): rice != 0 selects Rice */
int32_t ntc_column_code(const uint64_t* values, uint64_t n, int32_t rice, uint64_t* words, uint64_t cap, uint64_t* n_words,
                        uint64_t* param);

/* ---------------- file-level host pipeline ("next" row f3) ------------------------------------------
 * The driver loops of `ntcomp encode` (src/main.rs:141-181) and `ntcomp decode` (src/main.rs:183-211) as
 * overlapped stages: fastX parsing / block framing on a reader thread, the hot path on the GPU, the block
 * container (column codes + gzip, or gunzip + decode + zip) per 65,536-read block on n_threads host threads
 * (<= 0: all cores), and an in-order writer.  Paths may be "-" (stdin / stdout); fastX input may be
 * gzip-compressed.  encoded.dat layout: src/lib.rs:29-92,232-252.
 * encode stats[6]: reads, bases, records, bytes written, blocks written, blocks dropped (Q4);
 * decode stats[6]: sequences, FASTA bytes written, records, bytes read, blocks, 0.  stats may be NULL.
 * A damaged encoded.dat: ntc_decode_file writes every block that precedes the first damaged one, like the reference's
 * `while let Ok(..)` loop (src/main.rs:203), and then -- unlike the reference, which stops silently -- returns the error. */
int32_t ntc_encode_file(ntc_ctx* ctx, const char* fastx_path, const char* out_path, int32_t n_threads, uint64_t* stats);
int32_t ntc_decode_file(ntc_ctx* ctx, const char* in_path, const char* out_path, int32_t n_threads, uint64_t* stats);
/* Page-locked host memory for the buffers handed to the host-buffer entry points: copies to and from it run at full
 * PCIe rate without a staging pass (the file pipeline fills its read batches in such buffers).  Needs a GPU. */
int32_t ntc_pinned_alloc(uint64_t bytes, void** out);
void ntc_pinned_free(void* p);
/* Makes ctx's GPU the calling thread's current CUDA device.  Every entry point that takes a ctx does this itself; a helper
 * thread that only calls ntc_pinned_alloc / ntc_pinned_free has no ctx argument, and a new thread's current device is GPU 0,
 * so without this call its first allocation creates a CUDA context on GPU 0 (memory and start-up time there, failure where
 * GPU 0 is exclusive or hidden).  No counterpart in the reference (which has no devices). */
int32_t ntc_ctx_bind_thread(ntc_ctx* ctx);

/* fastX file (FASTA / FASTQ, plain or gzip) -> the normalised sequences (needletail normalize(true) for what can reach the
 * encoder, src/main.rs:158-163), parsed like the encode pipeline does: one reader thread cutting at record boundaries,
 * n_threads parsers.  *bases (n_bases bytes) and *offsets (n_reads + 1) are malloc'ed: release with ntc_free. */
int32_t ntc_fastx_load(const char* path, int32_t n_threads, uint8_t** bases, uint64_t** offsets, uint64_t* n_reads, uint64_t* n_bases);
void ntc_free(void* p);
/* `ntcomp build` front end (src/main.rs:97-134): every fastX record of every file is one input sequence */
int32_t ntc_index_build_from_files(const char* const* paths, uint64_t n_paths, uint32_t k, int32_t add_revcomp,
                                   int32_t n_threads, ntc_index** out);
/* the same with the construction on GPU `device` (ntc_index_build_gpu); n_threads parse the files */
int32_t ntc_index_build_from_files_gpu(int32_t device, const char* const* paths, uint64_t n_paths, uint32_t k, int32_t add_revcomp,
                                       int32_t n_threads, ntc_index** out);

/* ---------------- device-resident variants (buffers already in HBM, stream-ordered) ----------
 * Same semantics; all pointers are device pointers on ctx's GPU, `stream` is a cudaStream_t (0 =
 * default stream).  Results that are counts stay on the device: d_counts[0] = n_records / n_seqs,
 * d_counts[1] = bytes written.  ntc_ctx_status() synchronises the stream and returns the status of
 * everything enqueued since the last call (and the counts, if out != NULL). */
int32_t ntc_encode_device(ntc_ctx* ctx, const uint8_t* d_bases, const uint64_t* d_offsets, uint64_t n_reads,
                          uint64_t n_bases, uint64_t* d_records, uint64_t cap, uint64_t* d_rec_offsets,
                          uint64_t* d_counts, void* stream);
int32_t ntc_decode_device(ntc_ctx* ctx, const uint64_t* d_records, uint64_t n_records, uint64_t first_seq_number,
                          int32_t fasta, uint8_t* d_out, uint64_t cap, uint64_t* d_seq_offsets, uint64_t cap_seqs,
                          uint64_t* d_counts, void* stream);
int32_t ntc_ctx_status(ntc_ctx* ctx, void* stream);

/* ntc_decode_device for callers that must not allocate.  No counterpart in the reference: it exists for hosts of this
 * library, not for a function of ntcomp.  ntc_decode_device enlarges its workspaces with cudaFree + cudaMalloc whenever a
 * batch is bigger than any before it.  That is unacceptable to a host which has given all of HBM to its own memory pool
 * (the driver has nothing left to hand out, or the pool fragments), to a service that cannot afford the device-wide
 * synchronisation and driver lock of an allocation inside a request, and to a sandboxed plugin that may launch kernels
 * but not call the runtime's allocation functions.  This variant runs the same kernels and calls neither cudaMalloc nor
 * cudaFree nor any routine that does: its workspaces are carved from one array that is part of the library image, so
 * every GPU on which the library is loaded carries it (about 16 MiB) for the life of the process.
 *   - Limit: n_records <= NTC_STATIC_MAX_RECORDS (one 65,536-read block of short reads is ~0.3 M records); more returns
 *     NTC_ERR_CAPACITY with a message and enqueues nothing.  Cut a larger batch at sequence boundaries (just before a
 *     record whose `first` flag is set) and advance first_seq_number by the sequences already decoded.
 *   - The array is shared by every context and every stream of a GPU, and keeps what the last call left in it.  The call
 *     clears the part that has to start at zero itself; the rest is overwritten before it is read.  Two calls whose GPU
 *     work overlaps (different streams, or different contexts on one GPU) race on it: order them.
 *   - The first call on each GPU looks the array's device address up (cudaGetSymbolAddress) and keeps it; make that first
 *     call during set-up, before any stream capture is begun.
 * Arguments, results, status reporting (ntc_ctx_status) and output bytes are those of ntc_decode_device. */
#define NTC_STATIC_MAX_RECORDS 524288
uint64_t ntc_decode_static_max_records(void); /* NTC_STATIC_MAX_RECORDS of the loaded library */
int32_t ntc_decode_device_static(ntc_ctx* ctx, const uint64_t* d_records, uint64_t n_records, uint64_t first_seq_number,
                                 int32_t fasta, uint8_t* d_out, uint64_t cap, uint64_t* d_seq_offsets, uint64_t cap_seqs,
                                 uint64_t* d_counts, void* stream);

/* number of kernels this library launched on ctx since creation (bench.py's gpu_launches) */
uint64_t ntc_ctx_launch_count(const ntc_ctx* ctx);

/* measurement support (bench.py): per-kernel-class CUDA-event timing on the launching stream.
 * classes: 0 pack2bit, 1 matching statistics, 2 record-offset scan, 3 emit, 4 decode scan,
 * 5 sequence offsets, 6 walk, 7 misc, 8 column encode, 9 column decode.  ntc_ctx_kernel_times synchronises the device. */
#define NTC_KERNEL_CLASSES 10
int32_t ntc_ctx_set_profiling(ntc_ctx* ctx, int32_t on);
int32_t ntc_ctx_kernel_times(ntc_ctx* ctx, double* ms /*[NTC_KERNEL_CLASSES]*/, uint64_t* launches /*[NTC_KERNEL_CLASSES]*/, int32_t reset);
/* L2 random 32-byte-sector micro-benchmark: G sectors/s from a buffer of the given size */
int32_t ntc_l2_random_bench(ntc_ctx* ctx, uint64_t buffer_bytes, uint32_t loads_per_thread, int32_t dependent,
                            uint32_t n_blocks, double* gsectors_per_s);

/* matching statistics only (diagnostic / parity entry point for StreamingIndex::matching_statistics,
 * src/lib.rs:172-173): per base d (u8) and interval start (u32).  HOST buffers. */
int32_t ntc_matching_statistics(ntc_ctx* ctx, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads,
                                uint8_t* d_out, uint32_t* start_out);

/* ---------------- synthetic inputs (splitmix64 counter-based) -------------------- */
void ntc_synth_genome(uint64_t seed, uint64_t n, uint8_t* out);
/* reads r0..r0+n_reads of length read_len drawn from genome (both strands, substitution errors at
 * rate err_ppm / 1e6) */
void ntc_synth_reads(const uint8_t* genome, uint64_t genome_len, uint64_t seed, uint64_t r0, uint64_t n_reads,
                     uint32_t read_len, uint32_t err_ppm, uint8_t* out, int32_t n_threads);
/* same generator on the GPU, writing into device memory */
int32_t ntc_synth_reads_device(ntc_ctx* ctx, const uint8_t* d_genome, uint64_t genome_len, uint64_t seed, uint64_t r0,
                               uint64_t n_reads, uint32_t read_len, uint32_t err_ppm, uint8_t* d_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NTCOMP_B200_H */
