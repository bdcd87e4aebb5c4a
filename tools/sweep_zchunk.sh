mkdir -p gpurun_out
run() { # workload tag params...
  w=$1; tag=$2; shift 2
  timeout 200 python bench.py --workload $w --no-cpu --steps 10 --warmup 3 "$@" > gpurun_out/sw_${w}_$tag.json 2> gpurun_out/sw_${w}_$tag.err
  python -c "
import json
d=json.loads(open('gpurun_out/sw_${w}_$tag.json').read().strip().splitlines()[-1])
print('$w', '$tag', round(d['ms_per_step'],4), {k:round(x['ms_per_launch'],4) for k,x in d['kernels'].items() if 'stage' in k or 'smag' in k})
"
}
for z in 32 64 128 256; do run tgv3d zc$z --param zchunk=$z; done
for z in 8 12 16 24 32 64; do run tgv3d_f64 zc$z --param zchunk=$z; done
for z in 64 128 256; do run dt3d_smag szc$z --param smag_zchunk=$z; done
