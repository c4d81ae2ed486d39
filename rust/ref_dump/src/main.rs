// ref_dump <dir>: for every fixture <dir>/in/<name>.{k,contigs.fa,reads.fa} run the unmodified reference and write
//   <dir>/<name>.index.bin    u64 k, u64 n_sets, 4 x ceil(n/64) u64 row words (A,C,G,T; LSB first), n LCS bytes
//   <dir>/<name>.ms0.bin      matching statistics of read 0: per base u64 d, u64 start, u64 end   (lib.rs:172-173)
//   <dir>/<name>.records.bin  u64 count per read (n_reads words) followed by all record words     (lib.rs:163 + encode.rs:129)
//   <dir>/<name>.encoded.dat  file header + blocks of 65,536 reads, exactly as `ntcomp encode`    (main.rs:141-181)
//   <dir>/<name>.decoded.fa   `ntcomp decode` of that file                                        (main.rs:183-211)
// Accessor names of the sbwt crate (n_sets, sbwt().set_contains, LcsArray::access) are from its public API as recalled.
use std::io::{Read, Write};

use needletail::Sequence;
use sbwt::{SbwtIndexVariant, StreamingIndex};

fn read_fasta(path: &str) -> Vec<Vec<u8>> {
    let mut reader = needletail::parse_fastx_file(path).unwrap_or_else(|_| panic!("cannot read {}", path));
    let mut out = Vec::new();
    while let Some(rec) = reader.next() {
        out.push(rec.expect("fastX record").normalize(true).to_vec()); // main.rs:42, :163
    }
    out
}

fn put_u64(f: &mut std::fs::File, v: &[u64]) {
    for x in v {
        f.write_all(&x.to_le_bytes()).unwrap();
    }
}

fn main() {
    let dir = std::env::args().nth(1).expect("usage: ref_dump <dir>");
    let mut names: Vec<String> = std::fs::read_dir(format!("{}/in", dir)).expect("<dir>/in")
        .filter_map(|e| e.ok().and_then(|e| e.file_name().to_str().map(|s| s.to_string())))
        .filter_map(|s| s.strip_suffix(".k").map(|s| s.to_string())).collect();
    names.sort();
    for name in names {
        let k: usize = std::fs::read_to_string(format!("{}/in/{}.k", dir, name)).unwrap().trim().parse().unwrap();
        let contigs = read_fasta(&format!("{}/in/{}.contigs.fa", dir, name));
        let reads = read_fasta(&format!("{}/in/{}.reads.fa", dir, name));

        let mut opts = kbo::BuildOpts::default(); // as tests/fasta_data.rs:49-54 and main.rs:104-112
        opts.k = k;
        opts.dedup_batches = false;
        opts.temp_dir = None;
        opts.add_revcomp = true;
        opts.build_select = true;
        let (sbwt, lcs) = kbo::build(&contigs, opts);

        // ---- index arrays ----
        {
            let SbwtIndexVariant::SubsetMatrix(ix) = &sbwt;
            let n = ix.n_sets();
            let mut f = std::fs::File::create(format!("{}/{}.index.bin", dir, name)).unwrap();
            put_u64(&mut f, &[k as u64, n as u64]);
            for c in 0..4u8 {
                let mut row = vec![0u64; (n + 63) / 64];
                for i in 0..n {
                    if ix.sbwt().set_contains(i, c) {
                        row[i / 64] |= 1u64 << (i % 64);
                    }
                }
                put_u64(&mut f, &row);
            }
            let bytes: Vec<u8> = (0..n).map(|i| lcs.access(i) as u8).collect();
            f.write_all(&bytes).unwrap();

            // ---- matching statistics of read 0 ----
            let ms = StreamingIndex::new(ix, &lcs).matching_statistics(&reads[0]);
            let mut f = std::fs::File::create(format!("{}/{}.ms0.bin", dir, name)).unwrap();
            for (d, r) in ms {
                put_u64(&mut f, &[d as u64, r.start as u64, r.end as u64]);
            }
        }

        // ---- record stream per read ----
        let mut counts: Vec<u64> = Vec::new();
        let mut words: Vec<u64> = Vec::new();
        for r in &reads {
            let dict = ntcomp::encode_sequence(r, &sbwt, &lcs).unwrap();
            let rec = ntcomp::encode::encode_dictionary(&dict, &sbwt).unwrap();
            counts.push(rec.len() as u64);
            words.extend(rec);
        }
        let mut f = std::fs::File::create(format!("{}/{}.records.bin", dir, name)).unwrap();
        put_u64(&mut f, &counts);
        put_u64(&mut f, &words);

        // ---- encoded.dat: the loop of main.rs:152-177 ----
        let block_size = 65536;
        let mut dat: Vec<u8> = ntcomp::encode_file_header(0, 0, 0, 0).unwrap();
        let (mut at, mut r0) = (0usize, 0usize);
        while r0 < reads.len() {
            let r1 = (r0 + block_size).min(reads.len());
            let n_words: usize = counts[r0..r1].iter().sum::<u64>() as usize;
            let num = if r1 - r0 == block_size { block_size } else { reads.len() % block_size };
            let _ = ntcomp::write_block_to(&words[at..at + n_words], num, &mut dat); // errors dropped, as main.rs:170
            at += n_words;
            r0 = r1;
        }
        std::fs::write(format!("{}/{}.encoded.dat", dir, name), &dat).unwrap();

        // ---- decoded.fa: the loop of main.rs:191-209 ----
        let mut conn = std::io::Cursor::new(dat);
        let mut header = [0u8; 32];
        conn.read_exact(&mut header).unwrap();
        let file_header = ntcomp::decode_file_header(&header).unwrap();
        let mut out: Vec<u8> = Vec::new();
        let mut i = 0;
        while let Ok(records) = ntcomp::decode_block(&file_header, &sbwt, &mut conn) {
            for nucleotides in records {
                writeln!(&mut out, ">seq.{}", i + 1).unwrap();
                out.extend_from_slice(&nucleotides);
                out.push(b'\n');
                i += 1;
            }
        }
        std::fs::write(format!("{}/{}.decoded.fa", dir, name), &out).unwrap();
        eprintln!("{}: k = {}, {} contigs, {} reads, {} records", name, k, contigs.len(), reads.len(), words.len());
    }
}
