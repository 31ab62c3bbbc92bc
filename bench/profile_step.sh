mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/plain.log 2>&1 && DMPI_PROFILE_STEP=1 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_resnet50.csv python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/ncu1.log 2>&1
echo rc1=$?
python bench.py --model unet --steps 2 --warmup 3 --no-e2e > gpurun_out/plain2.log 2>&1 && DMPI_PROFILE_STEP=1 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_unet.csv python bench.py --model unet --steps 2 --warmup 3 --no-e2e > gpurun_out/ncu2.log 2>&1
echo rc2=$?
python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/plain3.log 2>&1 && DMPI_PROFILE_STEP=1 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:igemm_kernel -s 20 -c 4 -o gpurun_out/prof_igemm -f python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/ncu3.log 2>&1
echo rc3=$?
ls -la gpurun_out/*.ncu-rep
