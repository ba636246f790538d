/* integration/sam2bam.c — HARNESS: SAM text -> BAM + BAI with the reference's own I/O code
 * (src/sam.c, hts.c, bgzf.c; no samtools in this image).  SURVEY.md §7.1 step 2(ii), Appendix C.
 *   sam2bam in.sam out.bam      (writes out.bam and out.bam.bai; records must be sorted) */
#include <stdio.h>
#include "sam.h"

int main(int argc, char** argv) {
  if (argc != 3) { fprintf(stderr, "usage: sam2bam in.sam out.bam\n"); return 2; }
  samFile* in = sam_open(argv[1], "r", 0);
  if (!in) { fprintf(stderr, "cannot open %s\n", argv[1]); return 1; }
  bam_hdr_t* h = sam_hdr_read(in);
  samFile* out = sam_open(argv[2], "wb", 0);
  if (!h || !out) { fprintf(stderr, "cannot write %s\n", argv[2]); return 1; }
  sam_hdr_write(out, h);
  bam1_t* b = bam_init1();
  long n = 0;
  while (sam_read1(in, h, b) >= 0) { sam_write1(out, h, b); ++n; }
  bam_destroy1(b);
  bam_hdr_destroy(h);
  sam_close(in);
  sam_close(out);
  if (bam_index_build(argv[2], 0) != 0) { fprintf(stderr, "indexing failed\n"); return 1; }
  fprintf(stderr, "sam2bam: %ld records\n", n);
  return 0;
}
