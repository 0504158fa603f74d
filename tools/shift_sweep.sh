# on the GPU box: front-kernel batch size for long reads (BRB_FRONT_SHIFT pins log2 of the alignments per ticket; unset = measured per context)
for w in c3 c4; do for s in auto 3 0; do if [ $s = auto ]; then unset BRB_FRONT_SHIFT; else export BRB_FRONT_SHIFT=$s; fi; python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-workloads 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); ph=d['roofline']['phase_ms_per_step']; print('$w shift $s', round(d['ms_per_step'],4), [round(v,4) for v in ph.values() if v>0.001])"; done; done
