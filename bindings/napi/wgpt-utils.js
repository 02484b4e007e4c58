// Drop-in for the reference's webgpu-utils.js (class WebGPU) on top of the N-API addon, so that renderer.js runs
// unchanged apart from its import: same constructor / init(), same buffer helpers returning {data, buffer}, same
// device.createBindGroup / device.queue.writeBuffer / computePass / renderPass call shapes
// (/root/reference/renderer.js:12-16, 68-139, 163-191; /root/reference/webgpu-utils.js:15-55, 125-148).
// SOURCE ONLY on this box: there is no node here, so this file has never been executed.
//
//   webgpu-utils.js                                   here
//   init(): adapter + device + canvas context         nothing yet: the context needs W, H (first uniform write) and the
//                                                     WGSL constants (createShaderModule), so it is created lazily
//   createUniformBuffer / createStorageBuffer_*       {data, buffer: handle}; the handle remembers its typed array
//   createShaderModule(code)                          reads NUM_SAMPLES, MAX_BOUNCES, STRATIFY, IMPORTANCE_SAMPLING,
//                                                     STACK_SIZE out of the header text (header.wgsl:9-13): they are
//                                                     run-time fields of wgpt_config
//   device.createBindGroup({entries})                 gives every handle its binding number (renderer.js:109-123,130-136)
//   device.queue.writeBuffer(handle, 0, array)        addon.writeBuffer(ctx, binding, array)
//   computePass(pipeline, group, n)                   addon.dispatch(ctx, trunc(n + 1))   (webgpu-utils.js:130)
//   renderPass(...)                                   addon.present(ctx, canvasBytes); onPresent(canvasBytes, W, H)
import { createRequire } from 'node:module';
const addon = createRequire(import.meta.url)('./build/Release/wgpt.node');

const WGSL_CONSTANTS = {
  NUM_SAMPLES: 'numSamples', MAX_BOUNCES: 'maxBounces', STRATIFY: 'stratify',
  IMPORTANCE_SAMPLING: 'importanceSampling', STACK_SIZE: 'stackSize',
};

export class WebGPU {
  constructor() {
    this.ctx = null;
    this.options = { mode: 1 };          // WGPT_MODE_FAST; {mode: 0} selects the exact kernel
    this.handles = [];
    this.uniformHandle = null;
    this.onPresent = null;               // (Uint8ClampedArray rgba8, width, height) => void, e.g. write a PNG
    const self = this;
    this.device = {
      queue: { writeBuffer(handle, _offset, data) { self.#write(handle, data); } },
      createBindGroup(desc) {
        for (const e of desc.entries) e.resource.buffer.binding = e.binding;
        return { entries: desc.entries };
      },
    };
  }
  async init() { return this; }

  #handle(data, binding) {
    const h = { data, binding, dirty: true };
    this.handles.push(h);
    return { data, buffer: h };
  }
  createUniformBuffer(_label, data) { const r = this.#handle(data, 0); this.uniformHandle = r.buffer; return r; }
  createStorageBuffer_WriteOnly(_label, data) { return this.#handle(data, -1); }
  createStorageBuffer_ReadWrite(_label, data) { return this.#handle(data, -1); }
  createVertexBuffer(_label, data) { return { data, buffer: {} }; }     // raster-only (vertex.js): not needed
  createShaderModule(code) {
    for (const [name, field] of Object.entries(WGSL_CONSTANTS)) {
      const m = new RegExp('const\\s+' + name + '\\s*(?::\\s*\\w+)?\\s*=\\s*([\\w.]+)').exec(code);
      if (m) this.options[field] = m[1] === 'true' ? 1 : m[1] === 'false' ? 0 : Math.trunc(Number(m[1]));
    }
    return {};
  }
  createComputePipeline(_module) { return { getBindGroupLayout: () => ({}) }; }
  createRenderPipeline(_module) { return { getBindGroupLayout: () => ({}) }; }
  createRenderPassDescriptor() { return { colorAttachments: [{}] }; }

  #write(handle, data) {
    handle.data = data; handle.dirty = true;
    if (this.ctx && handle.binding >= 0) { addon.writeBuffer(this.ctx, handle.binding, data); handle.dirty = false; }
  }
  #ensureContext() {
    if (!this.ctx) {
      const u = this.uniformHandle.data;                     // Uniforms: screenDims first (header.wgsl:41-46)
      this.width = Math.trunc(u[0]); this.height = Math.trunc(u[1]);
      this.ctx = addon.create({ width: this.width, height: this.height, ...this.options });
      this.canvasBytes = new Uint8ClampedArray(4 * this.width * this.height);
    }
    for (const h of this.handles)
      if (h.dirty && h.binding >= 0 && h.data.length) { addon.writeBuffer(this.ctx, h.binding, h.data); h.dirty = false; }
  }
  computePass(_pipeline, _bindGroup, workGroupsNeeded) {
    this.#ensureContext();
    addon.dispatch(this.ctx, Math.trunc(workGroupsNeeded + 1));
  }
  renderPass(_descriptor, _pipeline, _bindGroup, _vertexBuffer) {
    this.#ensureContext();
    addon.present(this.ctx, this.canvasBytes);
    if (this.onPresent) this.onPresent(this.canvasBytes, this.width, this.height);
  }
  // headless extras (no counterpart in the page)
  render(firstFrame, nFrames) { this.#ensureContext(); addon.render(this.ctx, firstFrame, nFrames); }
  readFramebuffer() {
    this.#ensureContext();
    const out = new Float32Array(4 * this.width * this.height);
    addon.readBuffer(this.ctx, 3, out);
    return out;
  }
}
