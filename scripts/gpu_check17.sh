#!/bin/bash
# gpis_var_kernel variants side by side on one box: libcdx.so (padded-column MMA skip + record prefetch), libcdx_nopref.so
# (skip only), libcdx_noskip.so (neither = the previous kernel); parity suite on the first, C2 / C5-shape bench lines on all
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/c17_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/c17_pytest.log
for v in full nopref noskip; do
if [ $v != full ]; then export CDX_LIB=$PWD/compliancedex_b200/libcdx_$v.so; fi
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/c17_pg_$v.json 2> gpurun_out/c17_pg_$v.err; echo "bench exit $?"
for o in hammer coffeebottle; do
timeout 300 python bench.py --hand allegro --object $o --candidates 8192 --iters 100 --no-cpu-baseline > gpurun_out/c17_c5_${o}_$v.json 2> gpurun_out/c17_c5_${o}_$v.err; echo "bench exit $?"
done
done
unset CDX_LIB
timeout 600 python bench.py --workload gpis_c3 --queries 2000000 --var-mode -1 --no-cpu-baseline > gpurun_out/c17_c3_dmma.json 2> gpurun_out/c17_c3_dmma.err; echo "c3 exit $?"; cut -c1-300 gpurun_out/c17_c3_dmma.json
bash scripts/gpu_ncu4.sh
