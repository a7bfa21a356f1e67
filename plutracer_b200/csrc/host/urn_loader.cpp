// urn_loader.cpp — .urn scene loading (placeholder until the interpreter lands in this file).
#include "scene_store.h"

extern "C" int plu_scene_load_urn(const char*, const char* const*, int, plu_scene** out, char* err, int err_len) {
	if (out) *out = nullptr;
	pluhost::set_err(err, err_len, "urn loader not built into this libplugpu yet");
	return PLU_E_PARSE;
}
