python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc $?"; cut -c1-330 gpurun_out/bench.json
python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print('e2e',d['e2e']['value'],'roof',d['roofline']['frac'],d['roofline']['traffic'],'corr',d['roofline_corr']['frac'],d['roofline_corr']['levels'],'cpu',d['cpu_baseline']['value'], d['clocks'])"
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/fwd_launches.csv python tools/prof_forward.py > gpurun_out/prof_forward.log 2>&1
