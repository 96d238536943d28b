R=oracle/_ref
$R/ref_scene_dump $R/ref_scenes/room.json /tmp/room.scene.bin > /dev/null
python benchmarks/pt_stageprobe.py /tmp/room.scene.bin 1 2>&1 | tail -40
$R/ref_pathtracer $R/ref_scenes/room.json --iters 1 --out /tmp/ref1.bin > /dev/null
for gb in on off; do
cuda_pathtracer_b200/bin/cuda_path_tracer $R/ref_scenes/room.json --iters 1 --batch 1 --no-png --geom-bvh $gb --sum-out /tmp/got_$gb.bin > /dev/null
python - $gb <<'PY'
import numpy as np, sys
a=np.fromfile('/tmp/ref1.bin',dtype=np.float32).reshape(1200,1200,3); b=np.fromfile('/tmp/got_%s.bin'%sys.argv[1],dtype=np.float32).reshape(1200,1200,3)
ne=np.argwhere((a.view(np.uint32)!=b.view(np.uint32)).any(axis=2))
print(sys.argv[1],"differing pixels",len(ne),"of",1200*1200, "first", ne[:5].tolist(), "max abs", float(np.nanmax(np.abs(a-b))), "nonzero ref", int(np.count_nonzero(a)))
PY
done
