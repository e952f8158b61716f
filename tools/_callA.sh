set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/final_c3_n1.json 2> gpurun_out/final_c3_n1.err; tail -c 600 gpurun_out/final_c3_n1.json
python bench.py --sampler trilinear --no-cpu-baseline > gpurun_out/final_tri_n1.json 2> gpurun_out/final_tri_n1.err; tail -c 400 gpurun_out/final_tri_n1.json
python bench.py --steps 60 --warmup 3 --no-cpu-baseline > gpurun_out/pre_ncu.json 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/final_launches.csv python bench.py --steps 60 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo done
