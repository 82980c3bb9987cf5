# stage A time per pair vs batch size and frames per CTA, ws kernel vs the 4-byte-aligned batch kernel (no_align16=1)
one() { python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --no-ref-kernels --workload raft_realtime_640x480_b64_f16 "$@" | python -c "
import json,sys; d=json.loads(sys.stdin.read()); n=d['config']['pairs_per_gpu_per_step']; r=d['roofline']['stages']['rectify']; print('$*', 'us/pair=%.3f frac=%.3f' % (1e3*r['ms_per_launch']/n, r['frac']))"; }
for b in 64 128 256; do one --batch $b; one --batch $b --opt no_align16=1; done
for c in 2 4 8 16 32; do one --opt batch_ctas=$c; done
for c in 4 8 16; do one --opt batch_ctas=$c --opt no_align16=1; done
