smi() { nvidia-smi --query-gpu=clocks.sm,power.draw --format=csv,noheader,nounits -lms 100 > $1 & echo $!; }
[ -x tools/membench ] || nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/membench tools/membench.cu
P=$(smi gpurun_out/pw_xor.csv); ./tools/membench 3300 | tail -1; kill $P; sleep 3
P=$(smi gpurun_out/pw_fat.csv); python tools/sweep.py --variants 5 --reps 3000 | tail -1 | cut -c1-120; kill $P; sleep 3
P=$(smi gpurun_out/pw_fat4.csv); python tools/sweep.py --variants 5 --k 4 --p 2 --frag-mib 512 --reps 5000 | tail -1 | cut -c1-120; kill $P; sleep 3
P=$(smi gpurun_out/pw_thin.csv); python tools/sweep.py --variants 6 --reps 2500 | tail -1 | cut -c1-120; kill $P
for f in xor fat fat4 thin; do echo $f; tail -12 gpurun_out/pw_$f.csv | head -8 | tr '\n' ';'; echo; done
