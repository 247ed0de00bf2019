for v in 1 4 8 16 32; do echo "== vectors/thread $v"; DRN_BN_VPT=$v python bench/prof_bn2.py; done
