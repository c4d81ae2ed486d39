#!/usr/bin/env python
"""Regenerates rust/ntcomp-gpu.patch: the reference-side change that puts libntcomp_b200.so under `ntcomp encode` /
`ntcomp decode` behind a cargo feature `gpu` (INTEGRATION.md sections 3-4).  Needs a checkout of the reference
(read-only is fine: it is copied to a scratch directory first).
    python rust/make_patch.py /path/to/ntcomp > rust/ntcomp-gpu.patch
The patch adds src/gpu.rs (= rust/ntcomp-gpu/gpu.rs), a `gpu` feature + the -sys dependency in Cargo.toml, splits
decode_block into decode_block_records + decode_sequence (lib.rs:320-363) and adds the block-granular GPU loops to main.rs.
"""
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))


def sub(text, old, new):
    assert old in text, "reference changed: anchor not found:\n" + old
    return text.replace(old, new, 1)


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    tmp = tempfile.mkdtemp(prefix="ntc_patch_")
    for side in ("a", "b"):
        os.makedirs(os.path.join(tmp, side))
        shutil.copytree(os.path.join(ref, "src"), os.path.join(tmp, side, "src"))
        shutil.copy(os.path.join(ref, "Cargo.toml"), os.path.join(tmp, side, "Cargo.toml"))
    b = os.path.join(tmp, "b")
    shutil.copy(os.path.join(HERE, "ntcomp-gpu", "gpu.rs"), os.path.join(b, "src", "gpu.rs"))

    p = os.path.join(b, "Cargo.toml")
    s = open(p).read()
    s = sub(s, 'stderrlog = "0.6"\n',
            'stderrlog = "0.6"\n\n## B200 hot path (feature "gpu"): the FFI crate of ntcomp_b200 (rust/ntcomp-b200-sys there)\n'
            'ntcomp-b200-sys = { path = "../ntcomp_b200/rust/ntcomp-b200-sys", optional = true }\n\n[features]\ngpu = ["dep:ntcomp-b200-sys"]\n')
    open(p, "w").write(s)

    p = os.path.join(b, "src", "lib.rs")
    s = open(p).read()
    s = sub(s, "pub mod decode;\n", 'pub mod decode;\n#[cfg(feature = "gpu")]\npub mod gpu;\n')
    s = sub(s, "pub fn decode_block<R: std::io::Read>(\n    _file_header: &HeaderPlaceholder,\n    sbwt: &SbwtIndexVariant,\n    conn: &mut R,\n) -> Result<Vec<Vec<u8>>, E> {\n",
            "/// decode_block up to (not including) decode_sequence: the u64 records of the next block\n"
            "pub fn decode_block_records<R: std::io::Read>(\n    conn: &mut R,\n) -> Result<Vec<u64>, E> {\n")
    s = sub(s, "    let decoded = decode_sequence(&decompressed, sbwt);\n\n    Ok(decoded)\n}\n",
            "    Ok(decompressed)\n}\n\npub fn decode_block<R: std::io::Read>(\n    _file_header: &HeaderPlaceholder,\n    sbwt: &SbwtIndexVariant,\n"
            "    conn: &mut R,\n) -> Result<Vec<Vec<u8>>, E> {\n    let decompressed = decode_block_records(conn)?;\n\n"
            "    let decoded = decode_sequence(&decompressed, sbwt);\n\n    Ok(decoded)\n}\n")
    open(p, "w").write(s)

    p = os.path.join(b, "src", "main.rs")
    s = open(p).read()
    anchor = '            let mut reader = needletail::parse_fastx_file(query_file).unwrap_or_else(|_| panic!("Expected valid fastX file"));\n'
    s = sub(s, anchor, anchor + '''            #[cfg(feature = "gpu")]
            {
                // one GPU call per block of 65,536 reads: records == flat_map(encode_dictionary . encode_sequence)
                let gpu = ntcomp::gpu::Gpu::new(&sbwt, &lcs, 0).expect("B200 context");
                let mut bases: Vec<u8> = Vec::new();
                let mut offsets: Vec<u64> = vec![0];
                while let Some(rec) = read_from_fastx_parser(&mut *reader) {
                    bases.extend_from_slice(&rec.normalize(true));
                    offsets.push(bases.len() as u64);
                    if offsets.len() - 1 == block_size {
                        let u64_encoding = gpu.encode_reads(&bases, &offsets).unwrap();
                        let _ = ntcomp::write_block_to(&u64_encoding, block_size, &mut stdout);
                        bases.clear();
                        offsets.truncate(1);
                    }
                }
                if offsets.len() > 1 {
                    let u64_encoding = gpu.encode_reads(&bases, &offsets).unwrap();
                    let _ = ntcomp::write_block_to(&u64_encoding, offsets.len() - 1, &mut stdout);
                }
                let _ = stdout.flush();
                return;
            }
''')
    anchor = '            info!("Decoding encoded data...");\n'
    s = sub(s, anchor, anchor + '''            #[cfg(feature = "gpu")]
            {
                let (_, lcs) = kbo::index::load_sbwt(index_prefix.as_ref().unwrap());
                let gpu = ntcomp::gpu::Gpu::new(&sbwt, &lcs, 0).expect("B200 context");
                let mut i: u64 = 0;
                while let Ok(records) = ntcomp::decode_block_records(&mut conn) {
                    let (text, n_seqs) = gpu.decode_fasta(&records, i + 1).unwrap();
                    let _ = stdout.write_all(&text);
                    i += n_seqs;
                }
                let _ = stdout.flush();
                return;
            }
''')
    open(p, "w").write(s)
    out = subprocess.run(["diff", "-U2", "-rN", "a", "b"], cwd=tmp, capture_output=True, text=True).stdout
    lines = []
    for ln in out.splitlines():
        if ln.startswith("--- ") or ln.startswith("+++ "):
            ln = ln.split("\t")[0]                                       # no timestamps: the patch is reproducible
        lines.append(ln)
    sys.stdout.write("\n".join(lines) + "\n")
    shutil.rmtree(tmp)


if __name__ == "__main__":
    main()
