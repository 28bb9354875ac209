// translation unit "core" of libvdp_b200 (see vdp_all.inl)
#define VDP_EMIT_CORE 1
#include "vdp_all.inl"
