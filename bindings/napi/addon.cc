// Node N-API addon over libwgpt.so: the binding the reference's JavaScript host would load instead of
// talking to WebGPU.  It maps the typed arrays scene.js already produces
// (/root/reference/lib/scene.js: get_spheres() ... get_bvh(), Float32Array / Int32Array) to
// wgpt_write_buffer, computePass to wgpt_dispatch and renderPass to wgpt_present; wgpt-utils.js next to
// this file gives renderer.js the WebGPU class of webgpu-utils.js on top of it.
//
// SOURCE ONLY on this box: node and node_api.h are absent (probed; nothing can be installed), so this
// file is compiled nowhere here except for a syntax pass against a stand-in header that declares the
// N-API calls used below (tests/napi_stub/node_api.h, tests/test_host_and_abi.py).  Build where node
// exists: node-gyp rebuild (binding.gyp), with libwgpt.so on the loader path.
#include <node_api.h>
#include <stdint.h>
#include <string.h>
#include "wgpt.h"

namespace {

napi_value fail(napi_env env, const wgpt_ctx* ctx, const char* what) {
  const char* msg = wgpt_last_error(ctx);
  napi_throw_error(env, what, msg && *msg ? msg : what);
  return nullptr;
}

// reads an optional integer / boolean field of a plain JS object into *out
void field_i32(napi_env env, napi_value obj, const char* name, int32_t* out) {
  bool has = false;
  if (napi_has_named_property(env, obj, name, &has) != napi_ok || !has) return;
  napi_value v; napi_valuetype t;
  if (napi_get_named_property(env, obj, name, &v) != napi_ok || napi_typeof(env, v, &t) != napi_ok) return;
  if (t == napi_boolean) { bool b = false; napi_get_value_bool(env, v, &b); *out = b ? 1 : 0; }
  else if (t == napi_number) napi_get_value_int32(env, v, out);
}

wgpt_ctx* ctx_of(napi_env env, napi_value v) {
  void* p = nullptr;
  if (napi_get_value_external(env, v, &p) != napi_ok) return nullptr;
  return static_cast<wgpt_ctx*>(p);
}

void finalize_ctx(napi_env, void* data, void*) { wgpt_destroy(static_cast<wgpt_ctx*>(data)); }

// create({width, height, mode, device, numSamples, maxBounces, stratify, importanceSampling, stackSize,
//         tileRank, tileWorld, tileRows, strataRank, strataWorld}) -> context
// replaces WebGPU.init + pipeline creation (webgpu-utils.js:178-211, renderer.js:45-66,105-139); the WGSL
// constants of header.wgsl:9-13 arrive as fields.
napi_value Create(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  wgpt_config cfg; wgpt_default_config(&cfg);
  if (argc >= 1) {
    field_i32(env, argv[0], "width", &cfg.width);               field_i32(env, argv[0], "height", &cfg.height);
    field_i32(env, argv[0], "device", &cfg.device);             field_i32(env, argv[0], "mode", &cfg.mode);
    field_i32(env, argv[0], "numSamples", &cfg.num_samples);    field_i32(env, argv[0], "maxBounces", &cfg.max_bounces);
    field_i32(env, argv[0], "stratify", &cfg.stratify);         field_i32(env, argv[0], "importanceSampling", &cfg.importance_sampling);
    field_i32(env, argv[0], "stackSize", &cfg.stack_size);
    field_i32(env, argv[0], "tileRank", &cfg.tile_rank);        field_i32(env, argv[0], "tileWorld", &cfg.tile_world);
    field_i32(env, argv[0], "tileRows", &cfg.tile_rows);
    field_i32(env, argv[0], "strataRank", &cfg.strata_rank);    field_i32(env, argv[0], "strataWorld", &cfg.strata_world);
  }
  wgpt_ctx* ctx = nullptr;
  if (wgpt_create(&cfg, &ctx) != WGPT_OK) return fail(env, nullptr, "wgpt_create");
  napi_value ext;
  if (napi_create_external(env, ctx, finalize_ctx, nullptr, &ext) != napi_ok) { wgpt_destroy(ctx); return nullptr; }
  return ext;
}

// writeBuffer(ctx, binding, typedArray): createBuffer + queue.writeBuffer (webgpu-utils.js:15-55, renderer.js:184)
napi_value WriteBuffer(napi_env env, napi_callback_info info) {
  size_t argc = 3; napi_value argv[3];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  wgpt_ctx* ctx = argc >= 3 ? ctx_of(env, argv[0]) : nullptr;
  if (!ctx) { napi_throw_error(env, nullptr, "writeBuffer(ctx, binding, typedArray)"); return nullptr; }
  int32_t binding = -1; napi_get_value_int32(env, argv[1], &binding);
  napi_typedarray_type type; size_t length = 0, offset = 0; void* data = nullptr; napi_value ab;
  if (napi_get_typedarray_info(env, argv[2], &type, &length, &data, &ab, &offset) != napi_ok) {
    napi_throw_error(env, nullptr, "writeBuffer: third argument must be a typed array"); return nullptr;
  }
  size_t elem = 1;
  switch (type) {
    case napi_float32_array: case napi_int32_array: case napi_uint32_array: elem = 4; break;
    case napi_float64_array: elem = 8; break;
    case napi_int16_array: case napi_uint16_array: elem = 2; break;
    default: elem = 1; break;
  }
  if (wgpt_write_buffer(ctx, binding, data, length * elem) != WGPT_OK) return fail(env, ctx, "wgpt_write_buffer");
  return nullptr;
}

// dispatch(ctx, workgroups): computePass -> dispatchWorkgroups (webgpu-utils.js:125-134, renderer.js:187-188)
napi_value Dispatch(napi_env env, napi_callback_info info) {
  size_t argc = 2; napi_value argv[2];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  wgpt_ctx* ctx = argc >= 2 ? ctx_of(env, argv[0]) : nullptr;
  if (!ctx) { napi_throw_error(env, nullptr, "dispatch(ctx, workgroups)"); return nullptr; }
  uint32_t groups = 0; napi_get_value_uint32(env, argv[1], &groups);
  if (wgpt_dispatch(ctx, groups) != WGPT_OK) return fail(env, ctx, "wgpt_dispatch");
  return nullptr;
}

// render(ctx, firstFrame, nFrames): the batched renderAnimation loop (renderer.js:163-188, camera at rest)
napi_value Render(napi_env env, napi_callback_info info) {
  size_t argc = 3; napi_value argv[3];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  wgpt_ctx* ctx = argc >= 3 ? ctx_of(env, argv[0]) : nullptr;
  if (!ctx) { napi_throw_error(env, nullptr, "render(ctx, firstFrame, nFrames)"); return nullptr; }
  int32_t first = 1, n = 1; napi_get_value_int32(env, argv[1], &first); napi_get_value_int32(env, argv[2], &n);
  if (wgpt_render(ctx, first, n) != WGPT_OK) return fail(env, ctx, "wgpt_render");
  return nullptr;
}

// present(ctx, Uint8Array | Uint8ClampedArray rgba8 [, Float32Array rgba]): renderPass + fs (webgpu-utils.js:137-148,
// fragment.js:22-36); the canvas is the caller's array (e.g. the data of an ImageData)
napi_value Present(napi_env env, napi_callback_info info) {
  size_t argc = 3; napi_value argv[3];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  wgpt_ctx* ctx = argc >= 2 ? ctx_of(env, argv[0]) : nullptr;
  if (!ctx) { napi_throw_error(env, nullptr, "present(ctx, rgba8[, rgba])"); return nullptr; }
  napi_typedarray_type type; size_t length = 0, offset = 0; void* bytes = nullptr; void* floats = nullptr; napi_value ab;
  if (napi_get_typedarray_info(env, argv[1], &type, &length, &bytes, &ab, &offset) != napi_ok ||
      (type != napi_uint8_array && type != napi_uint8_clamped_array)) {
    napi_throw_error(env, nullptr, "present: rgba8 must be a Uint8Array or Uint8ClampedArray of 4*W*H bytes"); return nullptr;
  }
  if (argc >= 3) {
    size_t flen = 0;
    if (napi_get_typedarray_info(env, argv[2], &type, &flen, &floats, &ab, &offset) != napi_ok || type != napi_float32_array || flen != length) {
      napi_throw_error(env, nullptr, "present: rgba must be a Float32Array of 4*W*H floats"); return nullptr;
    }
  }
  if (wgpt_present(ctx, static_cast<float*>(floats), static_cast<uint8_t*>(bytes)) != WGPT_OK) return fail(env, ctx, "wgpt_present");
  return nullptr;
}

// readBuffer(ctx, binding, typedArray): headless read-back (no counterpart in the page)
napi_value ReadBuffer(napi_env env, napi_callback_info info) {
  size_t argc = 3; napi_value argv[3];
  napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
  wgpt_ctx* ctx = argc >= 3 ? ctx_of(env, argv[0]) : nullptr;
  if (!ctx) { napi_throw_error(env, nullptr, "readBuffer(ctx, binding, typedArray)"); return nullptr; }
  int32_t binding = -1; napi_get_value_int32(env, argv[1], &binding);
  napi_typedarray_type type; size_t length = 0, offset = 0; void* data = nullptr; napi_value ab;
  if (napi_get_typedarray_info(env, argv[2], &type, &length, &data, &ab, &offset) != napi_ok ||
      (type != napi_float32_array && type != napi_int32_array)) {
    napi_throw_error(env, nullptr, "readBuffer: third argument must be a Float32Array or Int32Array"); return nullptr;
  }
  if (wgpt_read_buffer(ctx, binding, data, length * 4) != WGPT_OK) return fail(env, ctx, "wgpt_read_buffer");
  return nullptr;
}

napi_value Init(napi_env env, napi_value exports) {
  struct { const char* name; napi_callback fn; } table[] = {
    {"create", Create}, {"writeBuffer", WriteBuffer}, {"dispatch", Dispatch}, {"render", Render},
    {"present", Present}, {"readBuffer", ReadBuffer},
  };
  for (const auto& e : table) {
    napi_value f;
    if (napi_create_function(env, e.name, NAPI_AUTO_LENGTH, e.fn, nullptr, &f) != napi_ok) return nullptr;
    napi_set_named_property(env, exports, e.name, f);
  }
  return exports;
}

}  // namespace

NAPI_MODULE(wgpt, Init)
