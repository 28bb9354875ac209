// translation unit "cvi" of libvdp_b200 (see vdp_all.inl)
#define VDP_EMIT_CVI 1
#include "vdp_all.inl"
