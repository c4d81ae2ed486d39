// src/gpu.rs -- ntcomp's hot path on a B200 through libntcomp_b200.so (feature "gpu").
//
// Same names and signatures as the CPU functions of lib.rs, with `&Gpu` in the place of the SBWT / LCS arguments:
//   encode_sequence  (lib.rs:163)   -> gpu::encode_sequence
//   encode_dictionary (encode.rs:129) is fused into gpu::Gpu::encode_reads (records come back packed)
//   decode_sequence  (lib.rs:254)   -> gpu::decode_sequence
// plus the two block-granular calls main.rs uses (one GPU call per 65,536-read block or more).
// sbwt-crate accessors used by `flatten` (k, n_sets, sbwt().set_contains, LcsArray::access) are written from the crate's
// public API as recalled; confirm against sbwt 0.3.11 where cargo exists.
use std::ops::Range;

use ntcomp_b200_sys as sys;
use sbwt::{LcsArray, SbwtIndexVariant};

type E = Box<dyn std::error::Error>;

pub struct Gpu {
    ctx: *mut sys::ntc_ctx,
}

impl Drop for Gpu {
    fn drop(&mut self) {
        unsafe { sys::ntc_ctx_destroy(self.ctx) }
    }
}

fn check(rc: i32) -> Result<(), E> {
    match rc {
        sys::NTC_OK => Ok(()),
        sys::NTC_ERR_EMPTY_READ => Err(Box::new(crate::encode::EncodeError {})), // lib.rs:224, encode.rs:133
        _ => Err(format!("libntcomp_b200 status {}: {}", rc, sys::last_error()).into()),
    }
}

impl Gpu {
    /// Flattens the loaded index (after kbo::index::load_sbwt, main.rs:149,190) and puts it on `device`.
    pub fn new(sbwt: &SbwtIndexVariant, lcs: &LcsArray, device: i32) -> Result<Gpu, E> {
        let SbwtIndexVariant::SubsetMatrix(sbwt) = sbwt;
        let n = sbwt.n_sets();
        let words = (n + 63) / 64;
        // rows: bit i of row c == node i has an outgoing edge labelled c (SubsetMatrix), LSB-first u64 words
        let mut rows = vec![vec![0u64; words]; 4];
        for i in 0..n {
            for c in 0..4 {
                if sbwt.sbwt().set_contains(i, c as u8) {
                    rows[c][i / 64] |= 1u64 << (i % 64);
                }
            }
        }
        let lcs_bytes: Vec<u8> = (0..n).map(|i| lcs.access(i) as u8).collect(); // k <= 255
        let mut idx = std::ptr::null_mut();
        check(unsafe {
            sys::ntc_index_from_arrays(sbwt.k() as u32, n as u64, rows[0].as_ptr(), rows[1].as_ptr(), rows[2].as_ptr(),
                                       rows[3].as_ptr(), lcs_bytes.as_ptr(), &mut idx)
        })?;
        let mut ctx = std::ptr::null_mut();
        let rc = unsafe { sys::ntc_ctx_create(device, idx, &mut ctx) };
        unsafe { sys::ntc_index_free(idx) };
        check(rc)?;
        Ok(Gpu { ctx })
    }

    /// flat_map(encode_dictionary . encode_sequence) over reads given as concatenated bases + n + 1 offsets
    /// (main.rs:166,169): the packed u64 records of all reads, in order.
    pub fn encode_reads(&self, bases: &[u8], offsets: &[u64]) -> Result<Vec<u64>, E> {
        let n_reads = (offsets.len() - 1) as u64;
        let mut rec = vec![0u64; unsafe { sys::ntc_encode_bound(bases.len() as u64) } as usize];
        let mut n = 0u64;
        check(unsafe {
            sys::ntc_encode_batch(self.ctx, bases.as_ptr(), offsets.as_ptr(), n_reads, rec.as_mut_ptr(), rec.len() as u64, &mut n,
                                  std::ptr::null_mut())
        })?;
        rec.truncate(n as usize);
        Ok(rec)
    }

    /// decode_sequence + the writer loop of main.rs:201-209: ">seq.N\n<bases>\n" for every sequence of the record stream,
    /// numbered from `first_seq_number`; returns the text and the number of sequences.
    pub fn decode_fasta(&self, records: &[u64], first_seq_number: u64) -> Result<(Vec<u8>, u64), E> {
        let (mut n_bases, mut n_seqs, mut bytes) = (0u64, 0u64, 0u64);
        unsafe { sys::ntc_decode_bound(records.as_ptr(), records.len() as u64, first_seq_number, &mut n_bases, &mut n_seqs, &mut bytes) };
        let mut text = vec![0u8; bytes as usize];
        let (mut nb, mut ns) = (0u64, 0u64);
        check(unsafe {
            sys::ntc_decode_fasta(self.ctx, records.as_ptr(), records.len() as u64, first_seq_number, text.as_mut_ptr(), bytes, &mut nb,
                                  &mut ns)
        })?;
        text.truncate(nb as usize);
        Ok((text, ns))
    }
}

/// lib.rs:163 with the GPU in the place of (sbwt, lcs)
pub fn encode_sequence(nucleotides: &[u8], gpu: &Gpu) -> Result<Vec<(usize, Range<usize>)>, E> {
    let offsets = [0u64, nucleotides.len() as u64];
    let cap = nucleotides.len().max(1);
    let mut tri = vec![0u64; 3 * cap];
    let mut n = 0u64;
    check(unsafe {
        sys::ntc_encode_sequence_batch(gpu.ctx, nucleotides.as_ptr(), offsets.as_ptr(), 1, tri.as_mut_ptr(), cap as u64, &mut n,
                                       std::ptr::null_mut())
    })?;
    Ok((0..n as usize).map(|i| (tri[3 * i] as usize, tri[3 * i + 1] as usize..tri[3 * i + 2] as usize)).collect())
}

/// encode.rs:129 with the GPU in the place of the sbwt (literal bases are rebuilt from the index, as access_kmer does)
pub fn encode_dictionary(dictionary: &[(usize, Range<usize>)], gpu: &Gpu) -> Result<Vec<u64>, E> {
    let tri: Vec<u64> = dictionary.iter().flat_map(|(l, r)| [*l as u64, r.start as u64, r.end as u64]).collect();
    let mut rec = vec![0u64; dictionary.len()];
    check(unsafe { sys::ntc_encode_dictionary(gpu.ctx, tri.as_ptr(), dictionary.len() as u64, rec.as_mut_ptr()) })?;
    Ok(rec)
}

/// lib.rs:254 with the GPU in the place of the sbwt
pub fn decode_sequence(encoding: &[u64], gpu: &Gpu) -> Vec<Vec<u8>> {
    let (mut n_bases, mut n_seqs, mut bytes) = (0u64, 0u64, 0u64);
    unsafe { sys::ntc_decode_bound(encoding.as_ptr(), encoding.len() as u64, 1, &mut n_bases, &mut n_seqs, &mut bytes) };
    let mut bases = vec![0u8; n_bases as usize];
    let mut off = vec![0u64; n_seqs as usize + 1];
    let mut ns = 0u64;
    let rc = unsafe {
        sys::ntc_decode_batch(gpu.ctx, encoding.as_ptr(), encoding.len() as u64, bases.as_mut_ptr(), n_bases, off.as_mut_ptr(), n_seqs,
                              &mut ns)
    };
    assert_eq!(rc, 0, "{}", sys::last_error()); // the CPU function panics on bad input too (access_kmer out of range)
    (0..ns as usize).map(|i| bases[off[i] as usize..off[i + 1] as usize].to_vec()).collect()
}
