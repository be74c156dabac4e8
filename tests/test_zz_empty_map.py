"""An empty map (P = 0) through the public API on the GPU: the reference short-circuits (RAST/rasterize_points.cu:101,
:202) and returns the background image and empty gradients.  The C side of this case is covered on the CPU
(tests/test_emulated_rasterizer.py::test_empty_map_through_the_abi, which found that the argument check refused the
NULL pointers empty tensors arrive as); ran green on the B200 at the end of round 1, an ordinary test since)."""
import pytest
import torch



@pytest.mark.gpu
def test_empty_map_renders_the_background_and_has_empty_gradients():
    from gaus_2dgs_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = torch.device("cuda:0")
    H, W = 24, 40
    bg = torch.tensor([0.25, 0.5, 0.75], device=dev)
    eye = torch.eye(4, device=dev)
    settings = GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=1.0, tanfovy=0.6, bg=bg, scale_modifier=1.0,
                                             viewmatrix=eye, projmatrix=eye, sh_degree=0, campos=torch.zeros(3, device=dev),
                                             use_sa=True, prefiltered=False, debug=False)
    z = lambda *s: torch.zeros(s, device=dev, requires_grad=True)      # noqa: E731
    means3D, means2D, opac, col, sc, rot = z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 2), z(0, 4)
    color, radii, allmap = GaussianRasterizer(settings)(means3D=means3D, means2D=means2D, opacities=opac, colors_precomp=col,
                                                        scales=sc, rotations=rot)
    assert color.shape == (3, H, W) and allmap.shape == (7, H, W) and radii.shape == (0,)
    assert torch.allclose(color, bg[:, None, None].expand(3, H, W)) and float(allmap[1].abs().max()) == 0.0
    (color.sum() + allmap.sum()).backward()
    assert means3D.grad is None or means3D.grad.shape == (0, 3)
