# stage A time per pair vs frames per CTA (batch_ctas = CTAs per SM the chunking aims at) and taper, ws kernel
one() { python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --no-ref-kernels --workload ${WL:-raft_realtime_640x480_b64_f16} "$@" | python -c "
import json,sys; d=json.loads(sys.stdin.read()); n=d['config']['pairs_per_gpu_per_step']; r=d['roofline']['stages']['rectify']; print('${WL:-raft} $*', 'us/pair=%.3f frac=%.3f' % (1e3*r['ms_per_launch']/n, r['frac']))"; }
one
for c in 8 12 16 20 24 32 48 64; do one --opt batch_ctas=$c; done
for t in 0 25 75; do one --opt chunk_taper=$t; done
for p in 0 1 4; do one --opt l2_prefetch=$p; done
WL=stream_1080p_b64_f32
one
for c in 8 16 32 64; do one --opt batch_ctas=$c; done
