for n in prof profhint; do echo "=== $n"; BFK_LIB=build/libbfk_$n.so timeout 120 python tools/gpu_chase_prof.py 20 1000 2>&1 | grep -v Warning | tail -16; done
for n in hint500 hint20k hint1m; do echo "=== $n"; BFK_LIB=build/libbfk_$n.so timeout 120 python tools/gpu_variant_check.py 20 1000 2>&1 | tail -1; done
