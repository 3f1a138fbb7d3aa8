set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
for a in "u32 15 8 256" "u32 16 8 128" "u32 17 4 128" "f64 14 8 256" "f64 15 8 128" "f64 16 4 128" "u64 14 8 256" "u64 15 8 128" "u64 16 4 128"; do
  timeout 120 lab/ntt_cluster_lab $a
done
timeout 120 lab/pipe_lab
