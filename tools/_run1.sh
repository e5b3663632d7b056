for i in 1 2; do
python tools/_ab/ab_chain.py tools/_ab/libvaex_old2.so 2>&1 | tail -n 1
python tools/_ab/ab_chain.py tools/_ab/libvaex_old.so 2>&1 | tail -n 1
python tools/_ab/ab_chain.py - 2>&1 | tail -n 1
done
