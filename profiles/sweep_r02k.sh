# r02k: occupancy variants again, on the kernel with separating-face hints and per-step level lists
K="python profiles/kbench.py --lib libexp_k9occ.so"
$K --tag 64x4x2_r128
MDR_MIN_BLOCKS=3 $K --smem-contacts 44 --max-contacts 160 --tag 64x4x3_r80_c44
$K --wpb 5 --max-contacts 160 --tag 64x5x2_r96
$K --wpb 6 --smem-contacts 48 --max-contacts 160 --tag 64x6x2_r80
MDR_MIN_BLOCKS=2 $K --max-contacts 160 --tag 64x4x2_r128_mc160
