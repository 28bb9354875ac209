// translation unit "prior" of libvdp_b200 (see vdp_all.inl)
#define VDP_EMIT_PRIOR 1
#include "vdp_all.inl"
