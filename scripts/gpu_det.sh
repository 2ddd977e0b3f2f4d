cd /root/repo
export B2CUDA_DEBUG_CHECKSUMS=1
for i in 1 2 3 4 5; do B2CUDA_ISLAND_MAX_BODIES=0 python scripts/gpu_determinism.py mixed 60 2 2>&1 | grep -E "first diff|checksum|nondet|debug:"; done
echo default; for i in 1 2 3 4 5; do python scripts/gpu_determinism.py mixed 60 2 2>&1 | grep -E "first diff|checksum|nondet|debug:"; done
