for ch in 37888 75776 151552 303104 606208; do echo "chunk $ch"; BFK_LW_CHUNK_ROWS=$ch B=606208 STEPS=5 timeout 300 python tools/gpu_lw_prof.py | tail -1; done
