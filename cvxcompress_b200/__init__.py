"""cvxcompress_b200 -- B200 (sm_100a) implementation of the CvxCompress block-wavelet
Compress / Decompress path.  The product is the shared library built from csrc/ (C ABI in
include/cvx_gpu.h, reference-compatible class in include/CvxCompress.hxx); this package is the
thin Python host side used by tests and bench.py."""
