for q in 0 2 3 4 5 8; do echo "VK_MMD_Q=$q"; VK_MMD_Q=$q timeout 100 python scripts/time_mmd.py 2>&1 | grep -E "type 0 passes [20]:"; done
