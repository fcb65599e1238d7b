"""B200-native post-processing for Leyan529/InstanceSegmentation- (YOLACT Detect + mask
assembly, Mask R-CNN NMS / RoIAlign / paste) behind the reference's own Python call
signatures.  Host side is plain Python + torch tensors for device memory and streams; all
compute is in libisb.so (hand-written CUDA for sm_100a, C ABI in include/isb.h).

There is no CPU path: every operator raises if the CUDA library is missing or a tensor is not
on a CUDA device.
"""
__version__ = "0.1.0"
