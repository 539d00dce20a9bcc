set -x
mkdir -p gpurun_out
timeout 300 python scripts/dbg_sn.py > gpurun_out/r02_c_dbg.log 2>&1; echo "dbg exit $?"; grep -v Warning gpurun_out/r02_c_dbg.log | tail -12
