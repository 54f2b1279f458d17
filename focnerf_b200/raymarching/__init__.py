from .raymarching import *  # noqa: F401,F403
from .raymarching import (compact_alive, march_rays_all, composite_rays_train_mse, composite_rays_train_mse_hc, WindowedMarch, composite_rays, composite_rays_train, march_rays, march_rays_train, morton3D,
                          morton3D_invert, near_far_from_aabb, packbits, sph_from_ray)
