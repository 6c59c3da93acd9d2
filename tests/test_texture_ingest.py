"""GPU: texture ingest on the device. The 8-bit level-0 images are converted and
mip-mapped in HBM (kernels.cuh: texture_level0_kernel / texture_mip_kernel); the pyramids must be
the host-built ones bit for bit — those are pinned to the reference through the image goldens —
and a frame rendered from them must be the same frame."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def textured(tmp_path_factory):
    from b200rt import scenes, standin
    from b200rt.host import HostScene
    d = tmp_path_factory.mktemp("tex")
    rng = np.random.default_rng(3)
    paths = []
    for name, (w, h) in (("square", (64, 64)), ("tall", (16, 64)), ("npot", (24, 20))):   # 256x1024-style non-square pyramid, non-power-of-two: no pyramid
        rgb = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        standin._write_tga(str(d / (name + ".tga")), rgb)
        paths.append(str(d / (name + ".tga")))
    scenes.write_sky(str(d / "sky.float"), 16)
    standin.mirror(str(d / "mirror.obj"), n=6)
    s = scenes.SceneWriter(160, 96)
    s.camera((0.0, 3.0, -8.0))
    s.ambient((0.3, 0.3, 0.3))
    s.sky(str(d / "sky.float"))
    s.plane((0.0, -1.0, 0.0), texture=paths[0])
    s.sphere(1.5, (-3.0, 1.0, 2.0), texture=paths[1])
    s.sphere(1.5, (3.0, 1.0, 2.0), texture=paths[2])
    s.mesh(str(d / "mirror.obj"), (0.0, 0.0, 6.0))
    s.point_light((60.0, 60.0, 60.0), (0.0, 8.0, -4.0))
    scene = HostScene(s.write(str(d / "tex.scene")))
    scene.update(0.0, 0)
    return scene.pack()


@pytest.mark.parametrize("filt", ["tri", "aniso", "ewa"])
def test_gpu_built_pyramids_and_frames_equal_host_built(textured, filt):
    from b200rt.device import Renderer
    pack = textured
    assert "s_texels_rgb8" in pack.a and pack.a["s_texels_rgb8"].size == 3 * (64 * 64 + 16 * 64 + 24 * 20)
    count = pack.a["s_texels"].size // 12
    host_texels = pack.a["s_texels"].view(np.float32).reshape(count, 3)

    a = Renderer(160, 96, 2, filt)
    a.upload_static(pack)
    frame_host = a.render(pack)
    assert np.array_equal(a.download_texels(0, count).view(np.uint32), host_texels.view(np.uint32))
    a.close()

    b = Renderer(160, 96, 2, filt)
    b.upload_static(pack, gpu_texture_ingest=True)
    frame_gpu = b.render(pack)
    assert np.array_equal(b.download_texels(0, count).view(np.uint32), host_texels.view(np.uint32))   # every level of every pyramid
    b.close()

    assert np.array_equal(frame_gpu["frame_buffer"], frame_host["frame_buffer"])
    assert np.array_equal(frame_gpu["presented"], frame_host["presented"])
