python -m pytest tests -q -x -m gpu 2>&1 | tail -3
python bench.py --no-box --no-cpu --variant samples 2>/dev/null | tail -1 > gpurun_out/bench_samples_n1.log
