timeout 300 python -m pytest tests/test_losses_frontend_gpu.py -x -q 2>&1 | tail -2
for cfg in "64 3000 1000 100" "64 1000 5000 300" "64 3000 2000 300"; do echo -n "cfg=$cfg : "; ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ctc_lattice -s 2 -c 1 --csv python tools/ctc_one.py $cfg 3 2>/dev/null | grep ctc_lattice | awk -F'","' '{print $NF}'; done
