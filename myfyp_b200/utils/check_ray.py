"""``ray_installed`` (reference: utils/check_ray.py:26-42).  Ray is never used here - simulation runs through
``learning/simulation.py`` (``FitPool``) - so this reports False; kept because scripts written for the reference call it."""


def ray_installed() -> bool:
    return False
