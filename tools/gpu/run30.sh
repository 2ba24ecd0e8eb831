set -x
cd $GRAFT_REPO_ROOT
(time python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')") > gpurun_out/r02zv_smoke.log 2>&1
tail -5 gpurun_out/r02zv_smoke.log
