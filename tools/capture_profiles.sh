# Captures of one round for profiles/: the bench line, the reference arm, the ncu launch list with DRAM bytes, --set full of the two serial kernels.
# Run on the GPU box from the repository root; everything lands in gpurun_out/ (scratch), tools/ncu_summ.py turns the CSVs into the summaries.
mkdir -p gpurun_out
set -x
cd /root/repo
timeout 900 python bench.py 2>gpurun_out/bench1_err.log | tail -1 > gpurun_out/r02_bench_1gpu_final.json
timeout 900 python bench.py --impl reference 2>gpurun_out/benchref_err.log | tail -1 > gpurun_out/r02_bench_reference_arm.json
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_ncu_launches_dram_512cube_final.csv python tools/prof_one.py 512 2 > gpurun_out/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:coeff_lattice -c 1 -o gpurun_out/lat_full -f python tools/prof_one.py 512 1 > gpurun_out/ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:huff_build -c 1 -o gpurun_out/huff_full -f python tools/prof_one.py 512 1 > gpurun_out/ncu3.log 2>&1
ncu -i gpurun_out/lat_full.ncu-rep --page raw --csv > gpurun_out/lat_full_raw.csv 2>/dev/null
ncu -i gpurun_out/huff_full.ncu-rep --page raw --csv > gpurun_out/huff_full_raw.csv 2>/dev/null
ls -la gpurun_out | tail -15
for f in gpurun_out/ncu1.log gpurun_out/ncu2.log gpurun_out/ncu3.log; do tail -n 2 "$f"; done
