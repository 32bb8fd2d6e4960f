"""bias_gan_b200 — B200-native hot path of Bias-GAN (reader -> stage -> G/D step).

Product code calls hand-written sm_100a kernels through libbgr.so (include/bgr.h).  There is no
CPU or library fallback; importing the kernels without the built library raises ImportError.
"""
__version__ = "0.1.0"
