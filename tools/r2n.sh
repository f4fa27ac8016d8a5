# rt_ply_ingest_static on the B200: the new test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2n_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/r2n_alone.log
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2n_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -6 gpurun_out/r2n_gpu_tests.log
