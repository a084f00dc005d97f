python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for ov in 0 1; do
HRC_TUNE_OVERLAP=$ov python bench.py --no-cpu-baseline --steps 16 > gpurun_out/sw_sg_ov$ov.json 2>/dev/null
HRC_TUNE_OVERLAP=$ov python bench.py --config material_grid --no-cpu-baseline --steps 16 > gpurun_out/sw_mg_ov$ov.json
HRC_TUNE_OVERLAP=$ov python bench.py --config cornell --no-cpu-baseline --steps 32 > gpurun_out/sw_cb_ov$ov.json
HRC_TUNE_OVERLAP=$ov python bench.py --config terrain --no-cpu-baseline --steps 8 > gpurun_out/sw_te_ov$ov.json
done
