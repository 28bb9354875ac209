// translation unit "vgp" of libvdp_b200 (see vdp_all.inl)
#define VDP_EMIT_VGP 1
#include "vdp_all.inl"
