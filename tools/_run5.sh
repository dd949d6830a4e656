python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for L in none "conv6_3.0,conv6_4.0,conv6a.0,conv6b.0" "conv6_0.0,conv6_1.0,conv6_2.0,conv6_3.0,conv6_4.0,conv6a.0,conv6b.0" "conv6_3.0,conv6_4.0,conv6a.0,conv6b.0,conv5_4.0"; do
  echo "== $L"; PWC_MMA_SMALL_LAYERS=$L python tools/layer_times.py 2>&1 | grep -E "graph replay"
  PWC_MMA_SMALL_LAYERS=$L python tools/layer_times.py 2>&1 | grep -E "graph replay"
done
