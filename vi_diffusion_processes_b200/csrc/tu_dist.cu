// translation unit "dist" of libvdp_b200 (see vdp_all.inl)
#define VDP_EMIT_DIST 1
#include "vdp_all.inl"
