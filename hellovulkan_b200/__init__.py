"""hellovulkan_b200 -- B200-native clustered-forward light culling (HelloVulkan's AABBGenerator.comp ->
LightCulling.comp pass pair) behind a C ABI (include/hvcf.h, hellovulkan_b200/lib/libhvcf.so)."""
from . import hvcf  # noqa: F401
