python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --no-cpu-baseline --steps 16 > gpurun_out/sw_sg.json 2>/dev/null
python bench.py --config material_grid --no-cpu-baseline --steps 16 > gpurun_out/sw_mg.json
python bench.py --config terrain --no-cpu-baseline --steps 8 > gpurun_out/sw_te.json
