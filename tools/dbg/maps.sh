for k in quad quarters half38 uniform1 uniform3 uniform0; do echo -n "$k: "; python tools/ablate.py 64 $k 1 | head -1; done
