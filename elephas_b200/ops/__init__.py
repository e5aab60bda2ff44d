"""Bindings of the hand-written sm_100a kernels (``csrc/``, ``libelephas_b200.so``) and the native
training / inference engine built from them."""
