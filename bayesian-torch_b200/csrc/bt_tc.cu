// placeholder until the tcgen05 path lands
#include "bt_common.cuh"
bool bt_tc_supported(const bt_geom*, const bt_layer_params*, bool, const char** why) { *why = "not built"; return false; }
int bt_tc_forward(const bt_geom*, const bt_layer_params*, const void*, void*, void*, bool, cudaStream_t) { bt_set_error("tc path not built"); return BT_ERR_UNSUPPORTED; }
int64_t bt_tc_workspace_bytes(const bt_geom*) { return 0; }
