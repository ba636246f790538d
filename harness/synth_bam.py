"""Synthetic BAM + BAI (+ FASTA, BED) for running the `otter` CLI (harness).

One chromosome; per region: 2 kb spacer, 1 kb left flank, a tandem repeat (allele 0 is the
reference), 1 kb right flank.  Reads span the region (flanks carry substitutions only, so the
CIGAR is `1000M <TR part> 1000M` with one I or D for the length difference of the repeat); a few
reads end inside the repeat (not spanning: they exercise invalid_reassignment).  The SAM text is
turned into BAM + BAI by integration/_build/sam2bam, i.e. by the reference's own src/sam.c."""
import os
import subprocess

import numpy as np

from . import synth

FLANK = 1000
SPACER = 2000


def _sub_only(rng, seq, e):
    out = seq.copy()
    hit = rng.random(len(seq)) < e
    out[hit] = synth.ALPHA[(synth._CODE[seq[hit]] + rng.integers(1, 4, int(hit.sum()))) % 4]
    return out


def write_dataset(outdir, n_regions, n_reads, len_lo, len_hi, e, seed=0, clipped_frac=0.1,
                  sam2bam=None):
    """-> dict(bam, bed, fasta, n_records).  Deterministic for a given seed."""
    os.makedirs(outdir, exist_ok=True)
    rng = np.random.default_rng([0xBA4, seed])
    chrom, beds, sam = [], [], []
    pos = 0
    for g in range(n_regions):
        tl = int(rng.integers(len_lo, len_hi + 1))
        mlen = int(rng.integers(20, 61)) if rng.random() < 0.1 else int(rng.integers(2, 7))
        motif = synth.rand_seq(rng, mlen)
        c0 = max(1, tl // mlen)
        c1 = c0 + int(rng.integers(0, int(c0 * 0.2) + 1))
        alleles = [np.tile(motif, c) for c in (c0, c1)]
        spacer, fl, fr = synth.rand_seq(rng, SPACER), synth.rand_seq(rng, FLANK), synth.rand_seq(rng, FLANK)
        chrom += [spacer, fl, alleles[0], fr]
        start = pos + SPACER + FLANK            # 0-based start of the repeat on the chromosome
        end = start + len(alleles[0])
        beds.append((start, end))
        lr = len(alleles[0])
        for r in range(n_reads):
            tr = synth.mutate(rng, alleles[int(rng.integers(0, 2))], e, homopolymer=e >= 0.02)
            left = _sub_only(rng, fl, e)
            if rng.random() < clipped_frac:      # ends inside the repeat: spanning on the left only
                keep = int(rng.integers(lr // 4, max(lr // 4 + 1, min(len(tr), lr) * 3 // 4)))
                seq = np.concatenate([left, tr[:keep]])
                cigar = f"{FLANK + keep}M"
            else:
                right = _sub_only(rng, fr, e)
                seq = np.concatenate([left, tr, right])
                lq = len(tr)
                mid = f"{min(lq, lr)}M" + (f"{lq - lr}I" if lq > lr else f"{lr - lq}D" if lr > lq else "")
                cigar = f"{FLANK}M{mid}{FLANK}M"
            sam.append(f"r{g}_{r}\t0\tchr1\t{start - FLANK + 1}\t60\t{cigar}\t*\t0\t0\t{seq.tobytes().decode()}\t*\trq:f:0.999")
        pos = end + FLANK
    chrom.append(synth.rand_seq(rng, SPACER))
    ref = np.concatenate(chrom)
    fasta = os.path.join(outdir, "ref.fa")
    with open(fasta, "w") as f:
        f.write(">chr1\n")
        s = ref.tobytes().decode()
        for i in range(0, len(s), 60):
            f.write(s[i:i + 60] + "\n")
    with open(fasta + ".fai", "w") as f:
        f.write(f"chr1\t{len(ref)}\t6\t60\t61\n")
    bed = os.path.join(outdir, "regions.bed")
    with open(bed, "w") as f:
        for a, b in beds:
            f.write(f"chr1\t{a}\t{b}\n")
    samp = os.path.join(outdir, "reads.sam")
    with open(samp, "w") as f:
        f.write(f"@HD\tVN:1.6\tSO:coordinate\n@SQ\tSN:chr1\tLN:{len(ref)}\n")
        f.write("\n".join(sam) + "\n")
    bam = os.path.join(outdir, "reads.bam")
    tool = sam2bam or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                   "integration", "_build", "sam2bam")
    subprocess.check_call([tool, samp, bam], stderr=subprocess.DEVNULL)
    os.remove(samp)
    return dict(bam=bam, bed=bed, fasta=fasta, n_records=len(sam))
