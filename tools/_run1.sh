set -x
timeout 900 python -m pytest tests -x -q -m gpu -k "host_path or golden or ref_" 2>&1 | tail -3
for nt in 0 1 0 1; do echo "NT $nt"; B200SORT_COPY_NT=$nt timeout 300 python tools/e2e_probe.py 5e8 2>&1 | grep -i "pageable"; done
