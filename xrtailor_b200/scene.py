"""Scene generators live in the top-level `workloads` package (they are inputs, not product code: the reference arm of
bench.py uses them without loading libxrt_b200.so).  This module re-exports them for the tests and adds make_solver,
the one helper that does touch the product."""
from workloads.scene import *  # noqa: F401,F403
from workloads.scene import SEED, Scene, _BODY_PARTS  # noqa: F401


def make_solver(scene, colors=None, bindings=None, lra=None, quality=None, strain_colors=None, **params):
    """Reference call order: AddCloth, AddObstacle, constraints, (colours), Start."""
    from .solver import ClothSolver
    if "particle_diameter" not in params:
        params["particle_diameter"] = scene.particle_diameter()
    s = ClothSolver(**params)
    s.AddCloth(scene.cloth_pos, scene.cloth_tri, scene.cloth_edges, scene.inv_mass)
    if scene.obs_pos is not None:
        s.AddObstacle(scene.obs_pos, scene.obs_tri, scene.obs_edges)
    s.SetStretch(scene.stretch_idx, scene.stretch_rest)
    s.SetBend(scene.bend_idx, scene.bend_theta)
    if bindings is not None:
        s.AddBinding(*bindings)
    if lra is not None:
        s.AddGeodesic(*lra)
    if colors is not None:
        s.SetColors(0, colors[0])
        s.SetColors(1, colors[1])
    if quality is not None:  # Quality-mode handle (quality: XrtQualityParams-like, True for the defaults)
        s.SetQuality(None if quality is True else quality, strain_colors)
    s.Start()
    return s
