//! Raw bindings to `libntcomp_b200.so`: one declaration per symbol of `include/ntcomp_b200.h`, same order, same
//! argument meaning (see the header for the reference file:line each entry point stands for).
//! `tests/test_abi.py::test_rust_sys_crate_declares_every_header_symbol` keeps this file and the header in step.
//! NOT compiled in the repository's own image (no Rust toolchain there): written against the header by hand.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_void};

#[repr(C)]
pub struct ntc_index {
    _private: [u8; 0],
}
#[repr(C)]
pub struct ntc_ctx {
    _private: [u8; 0],
}

pub const NTC_OK: i32 = 0;
pub const NTC_ERR_INVALID_ARG: i32 = 1;
pub const NTC_ERR_CUDA: i32 = 2;
pub const NTC_ERR_CAPACITY: i32 = 3;
pub const NTC_ERR_EMPTY_READ: i32 = 4;
pub const NTC_ERR_BAD_BASE: i32 = 5;
pub const NTC_ERR_LEN24: i32 = 6;
pub const NTC_ERR_IO: i32 = 7;
pub const NTC_ERR_BAD_RECORD: i32 = 8;
pub const NTC_ERR_UNSUPPORTED: i32 = 9;
pub const NTC_ERR_EMPTY_COLUMN: i32 = 10;

pub const NTC_USE_ENCODE: u32 = 1;
pub const NTC_USE_DECODE: u32 = 2;
pub const NTC_BLOCK_FULL_LAST_LITERAL_WORD: i32 = 1;
pub const NTC_KERNEL_CLASSES: usize = 10;

/// `ntc_coded_column`: one coded column of a block (BlockHeader fields, src/lib.rs:38-50)
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct ntc_coded_column {
    pub n_values: u64,
    pub n_words: u64,
    pub param: u64,
    pub word_offset: u64,
}
/// `ntc_coded_block`: one 65,536-read block as four coded columns
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct ntc_coded_block {
    pub n_records: u64,
    pub num_reads: u64,
    pub status: i32,
    pub reserved: i32,
    pub col: [ntc_coded_column; 4],
}

extern "C" {
    pub fn ntc_last_error() -> *const c_char;
    pub fn ntc_version() -> *const c_char;

    // ---- index (host) ----
    pub fn ntc_index_build(seqs: *const *const u8, lens: *const u64, n_seqs: u64, k: u32, add_revcomp: i32, n_threads: i32,
                           out: *mut *mut ntc_index) -> i32;
    /// `ntcomp build` on the GPU (row f4): same index as ntc_index_build; stats = 8 f64 or null
    pub fn ntc_index_build_gpu(device: i32, seqs: *const *const u8, lens: *const u64, n_seqs: u64, k: u32, add_revcomp: i32,
                               out: *mut *mut ntc_index, stats: *mut f64) -> i32;
    pub fn ntc_index_from_arrays(k: u32, n_sets: u64, row_a: *const u64, row_c: *const u64, row_g: *const u64, row_t: *const u64,
                                 lcs: *const u8, out: *mut *mut ntc_index) -> i32;
    pub fn ntc_index_save(idx: *const ntc_index, path: *const c_char) -> i32;
    pub fn ntc_index_load(path: *const c_char, out: *mut *mut ntc_index) -> i32;
    pub fn ntc_index_k(idx: *const ntc_index) -> u32;
    pub fn ntc_index_n_sets(idx: *const ntc_index) -> u64;
    pub fn ntc_index_get_C(idx: *const ntc_index, out: *mut u64);
    pub fn ntc_index_get_row(idx: *const ntc_index, c: i32, out: *mut u64);
    pub fn ntc_index_get_lcs(idx: *const ntc_index, out: *mut u8);
    pub fn ntc_index_free(idx: *mut ntc_index);

    // ---- context ----
    pub fn ntc_ctx_create(device: i32, idx: *const ntc_index, out: *mut *mut ntc_ctx) -> i32;
    pub fn ntc_ctx_create_for(device: i32, idx: *const ntc_index, uses: u32, out: *mut *mut ntc_ctx) -> i32;
    pub fn ntc_ctx_destroy(ctx: *mut ntc_ctx);
    pub fn ntc_ctx_index_bytes(ctx: *const ntc_ctx, out: *mut u64);
    pub fn ntc_ctx_layout_read(ctx: *mut ntc_ctx, which: i32, out: *mut c_void, cap: u64, n_bytes: *mut u64) -> i32;
    pub fn ntc_ctx_set_tile_len(ctx: *mut ntc_ctx, tile_len: u32) -> i32;
    pub fn ntc_ctx_set_copy_directions(ctx: *mut ntc_ctx, serial: i32) -> i32;
    pub fn ntc_ctx_set_host_threads(ctx: *mut ntc_ctx, n_threads: i32) -> i32;

    // ---- encode ----
    pub fn ntc_encode_batch(ctx: *mut ntc_ctx, bases: *const u8, offsets: *const u64, n_reads: u64, records: *mut u64, cap: u64,
                            n_records: *mut u64, rec_offsets: *mut u64) -> i32;
    pub fn ntc_encode_sequence_batch(ctx: *mut ntc_ctx, bases: *const u8, offsets: *const u64, n_reads: u64, triples: *mut u64,
                                     cap: u64, n_records: *mut u64, rec_offsets: *mut u64) -> i32;
    pub fn ntc_encode_dictionary(ctx: *mut ntc_ctx, triples: *const u64, n: u64, records: *mut u64) -> i32;
    pub fn ntc_encode_bound(n_bases: u64) -> u64;

    // ---- decode ----
    pub fn ntc_decode_batch(ctx: *mut ntc_ctx, records: *const u64, n_records: u64, bases: *mut u8, cap: u64, seq_offsets: *mut u64,
                            cap_seqs: u64, n_seqs: *mut u64) -> i32;
    pub fn ntc_decode_fasta(ctx: *mut ntc_ctx, records: *const u64, n_records: u64, first_seq_number: u64, text: *mut u8, cap: u64,
                            n_bytes: *mut u64, n_seqs: *mut u64) -> i32;
    pub fn ntc_decode_bound(records: *const u64, n_records: u64, first_seq_number: u64, n_bases: *mut u64, n_seqs: *mut u64,
                            fasta_bytes: *mut u64);

    // ---- block container ----
    pub fn ntc_block_bound(n_records: u64) -> u64;
    pub fn ntc_block_write(records: *const u64, n_records: u64, num_reads: u64, out: *mut u8, cap: u64, n_bytes: *mut u64) -> i32;
    pub fn ntc_block_read(data: *const u8, size: u64, records: *mut u64, cap: u64, n_records: *mut u64, consumed: *mut u64) -> i32;
    pub fn ntc_block_read_ex(data: *const u8, size: u64, records: *mut u64, cap: u64, n_records: *mut u64, consumed: *mut u64,
                             flags: i32) -> i32;
    pub fn ntc_block_write_gpu(ctx: *mut ntc_ctx, records: *const u64, n_records: u64, num_reads: u64, out: *mut u8, cap: u64,
                               n_bytes: *mut u64) -> i32;
    pub fn ntc_block_read_gpu(ctx: *mut ntc_ctx, data: *const u8, size: u64, records: *mut u64, cap: u64, n_records: *mut u64,
                              consumed: *mut u64, flags: i32) -> i32;

    // ---- fused, device-resident column codec ----
    pub fn ntc_encode_columns(ctx: *mut ntc_ctx, bases: *const u8, offsets: *const u64, n_reads: u64, block_reads: u64,
                              blocks: *mut ntc_coded_block, cap_blocks: u64, n_blocks: *mut u64, words: *mut *const u64,
                              n_words: *mut u64, n_records: *mut u64) -> i32;
    pub fn ntc_block_assemble(blk: *const ntc_coded_block, words: *const u64, out: *mut u8, cap: u64, n_bytes: *mut u64) -> i32;
    pub fn ntc_sub_block_assemble(blk: *const ntc_coded_block, column: i32, words: *const u64, out: *mut u8, cap: u64,
                                  n_bytes: *mut u64) -> i32;
    pub fn ntc_block_disassemble(data: *const u8, size: u64, blk: *mut ntc_coded_block, words: *mut u64, cap_words: u64,
                                 n_words: *mut u64, consumed: *mut u64) -> i32;
    pub fn ntc_decode_columns_fasta(ctx: *mut ntc_ctx, blocks: *const ntc_coded_block, n_blocks: u64, words: *const u64, n_words: u64,
                                    first_seq_number: u64, flags: i32, text: *mut *const u8, n_bytes: *mut u64, n_seqs: *mut u64,
                                    n_records: *mut u64) -> i32;
    pub fn ntc_column_code(values: *const u64, n: u64, rice: i32, words: *mut u64, cap: u64, n_words: *mut u64, param: *mut u64) -> i32;

    // ---- file-level host pipeline ----
    pub fn ntc_encode_file(ctx: *mut ntc_ctx, fastx_path: *const c_char, out_path: *const c_char, n_threads: i32, stats: *mut u64) -> i32;
    pub fn ntc_decode_file(ctx: *mut ntc_ctx, in_path: *const c_char, out_path: *const c_char, n_threads: i32, stats: *mut u64) -> i32;
    pub fn ntc_pinned_alloc(bytes: u64, out: *mut *mut c_void) -> i32;
    pub fn ntc_pinned_free(p: *mut c_void);
    pub fn ntc_ctx_bind_thread(ctx: *mut ntc_ctx) -> i32;
    pub fn ntc_fastx_load(path: *const c_char, n_threads: i32, bases: *mut *mut u8, offsets: *mut *mut u64, n_reads: *mut u64,
                          n_bases: *mut u64) -> i32;
    pub fn ntc_free(p: *mut c_void);
    pub fn ntc_index_build_from_files(paths: *const *const c_char, n_paths: u64, k: u32, add_revcomp: i32, n_threads: i32,
                                      out: *mut *mut ntc_index) -> i32;
    pub fn ntc_index_build_from_files_gpu(device: i32, paths: *const *const c_char, n_paths: u64, k: u32, add_revcomp: i32, n_threads: i32,
                                          out: *mut *mut ntc_index) -> i32;

    // ---- device-resident variants (device pointers, cudaStream_t) ----
    pub fn ntc_encode_device(ctx: *mut ntc_ctx, d_bases: *const u8, d_offsets: *const u64, n_reads: u64, n_bases: u64,
                             d_records: *mut u64, cap: u64, d_rec_offsets: *mut u64, d_counts: *mut u64, stream: *mut c_void) -> i32;
    pub fn ntc_decode_device(ctx: *mut ntc_ctx, d_records: *const u64, n_records: u64, first_seq_number: u64, fasta: i32,
                             d_out: *mut u8, cap: u64, d_seq_offsets: *mut u64, cap_seqs: u64, d_counts: *mut u64,
                             stream: *mut c_void) -> i32;
    pub fn ntc_ctx_status(ctx: *mut ntc_ctx, stream: *mut c_void) -> i32;
    // allocation-free decode on the library's static scratch (no counterpart in the reference; see the header)
    pub fn ntc_decode_static_max_records() -> u64;
    pub fn ntc_decode_device_static(ctx: *mut ntc_ctx, d_records: *const u64, n_records: u64, first_seq_number: u64, fasta: i32,
                                    d_out: *mut u8, cap: u64, d_seq_offsets: *mut u64, cap_seqs: u64, d_counts: *mut u64,
                                    stream: *mut c_void) -> i32;

    // ---- measurement support ----
    pub fn ntc_ctx_launch_count(ctx: *const ntc_ctx) -> u64;
    pub fn ntc_ctx_set_profiling(ctx: *mut ntc_ctx, on: i32) -> i32;
    pub fn ntc_ctx_kernel_times(ctx: *mut ntc_ctx, ms: *mut f64, launches: *mut u64, reset: i32) -> i32;
    pub fn ntc_l2_random_bench(ctx: *mut ntc_ctx, buffer_bytes: u64, loads_per_thread: u32, dependent: i32, n_blocks: u32,
                               gsectors_per_s: *mut f64) -> i32;
    pub fn ntc_matching_statistics(ctx: *mut ntc_ctx, bases: *const u8, offsets: *const u64, n_reads: u64, d_out: *mut u8,
                                   start_out: *mut u32) -> i32;

    // ---- synthetic inputs ----
    pub fn ntc_synth_genome(seed: u64, n: u64, out: *mut u8);
    pub fn ntc_synth_reads(genome: *const u8, genome_len: u64, seed: u64, r0: u64, n_reads: u64, read_len: u32, err_ppm: u32,
                           out: *mut u8, n_threads: i32);
    pub fn ntc_synth_reads_device(ctx: *mut ntc_ctx, d_genome: *const u8, genome_len: u64, seed: u64, r0: u64, n_reads: u64,
                                  read_len: u32, err_ppm: u32, d_out: *mut u8, stream: *mut c_void) -> i32;
}

/// the calling thread's last error message
pub fn last_error() -> String {
    unsafe { std::ffi::CStr::from_ptr(ntc_last_error()).to_string_lossy().into_owned() }
}
