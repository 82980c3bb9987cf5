for n in 1 2 3 7 64; do timeout 40 python tools/dbg_tma.py $n 2>&1 | tail -3; echo "rc=$?"; done
timeout 40 python tools/dbg_tma.py 5 tests/golden/calib_b.yml 2>&1 | tail -3
timeout 40 python tools/dbg_tma.py 5 tests/golden/calib_c_rational.yml 2>&1 | tail -3
