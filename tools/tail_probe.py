import sys
sys.path.insert(0,'cpu-raytracer_b200/python'); sys.path.insert(0,'.')
import bench
from b200rt.device import Renderer
sf, scene = bench.build_scene(); pack = scene.pack()
for stride in (1,2,4,8):
    r = Renderer(1920,1080,3,"tri",tile_stride=stride,tile_phase=0); r.upload_static(pack)
    for _ in range(3): s = r.render(pack, want_presented=False)["stats"]
    print(stride, "render ms", round(s["device_ms_render"],3), "tail ms", round(s["device_ms_tail"],3))
    r.close()
