"""genome_weaver_align_b200 — B200-native (sm_100a) single-end read alignment path of genome-weaver-align (BSF strategy).

The compute path is the CUDA library libgwalign.so (csrc/, C-ABI in include/gw_align.h); this package is the thin host-side
mirror of the reference's aligner interface. There is no CPU implementation behind it."""
from .aligner import (AlignmentBatch, AlignmentConfig, BidirectionalSuffixFilter, FMIndexOnGenome, SAMOutput, alignBlockDetailed, bwt,  # noqa: F401
                      cigar_string, cyclic_suffix_array, staircase_table)
