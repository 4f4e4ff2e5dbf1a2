"""visionaray_b200 -- B200-native (sm_100a) ray-query hot path behind the reference's API.

See DESIGN.md.  `_native` binds the in-tree shared libraries; `scenes` wraps the host C++ scene
generators and the binned-SAH builder; `api` is the Python mirror of the host adaptor
(scene upload, closest_hit / any_hit, frame()).
"""
__version__ = "0.1.0"
