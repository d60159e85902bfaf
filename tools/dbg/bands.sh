for b in 1 2 3 5 8 64; do echo bands=$b; TDLT_INV_BANDS=$b python tools/ablate.py 64 quad 1 | head -1; done
