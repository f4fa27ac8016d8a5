"""Small seeded scenes whose oracle renders are frozen under tests/golden/ (tools/make_golden.py)."""
from myraytracer_b200 import scenegen

GOLDEN_CASES = {
    "c1_small": (scenegen.config_c1, dict(width=96, height=96, samples=1)),
    "c2_small": (scenegen.config_c2, dict(n=40, width=96, height=96, samples=1)),
    "c3_small": (scenegen.config_c3, dict(width=64, height=64, samples=4, depth=4, ico_level=1)),
    "c5_small": (scenegen.config_c5, dict(n=24, width=96, height=54, samples=4, meshes=3)),
    # beyond BASELINE.json's configs: plane + MeshInstances; glossy materials + thin-lens camera
    "instances_small": (scenegen.config_instances, dict(width=64, height=48, samples=4)),
    "c3_glossy_dof_small": (scenegen.config_c3, dict(width=48, height=48, samples=4, depth=3, ico_level=1, roughness=0.2,
                                                     aperture=0.2, focus=14.0)),
    # area lights: the per-chunk jitter index (Object+Extension.swift:145-186,288)
    "c3_area_small": (scenegen.config_c3, dict(width=56, height=44, samples=4, depth=3, ico_level=1, area_lights=3)),
    # motion blur: Mesh.motionBlur (triangle test) and MeshInstance.motionBlur (static instance boxes: the reference's
    # own TLAS topology decides which displaced instances a ray reaches)
    "motion_small": (scenegen.config_instances, dict(width=64, height=48, samples=4, mesh_motion=(0.4, -0.3, 0.2),
                                                     instance_motion=1.5)),
}
