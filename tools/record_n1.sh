for c in 0 1 2 4; do
  python bench.py --config $c > gpurun_out/r02_scale_cfg${c}_n1.json 2> gpurun_out/r02_scale_cfg${c}_n1.err; echo "cfg $c rc=$?"
done
python bench.py --impl reference --config 0 --steps 3 --warmup 1 > gpurun_out/r02_reference_cfg0.json 2> gpurun_out/r02_reference_cfg0.err; echo "ref rc=$?"
tail -c 600 gpurun_out/r02_scale_cfg0_n1.json
