mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/t.log
python bench.py --impl reference > gpurun_out/bref_default.json 2> gpurun_out/bref_default.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/bref_default.json
python bench.py > gpurun_out/b_default.json 2> gpurun_out/b_default.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/b_default.json
