{
  "targets": [{
    "target_name": "wgpt",
    "sources": ["addon.cc"],
    "include_dirs": ["../../include"],
    "libraries": ["-L<(module_root_dir)/../../webgpu_path_tracer_b200", "-lwgpt",
                  "-Wl,-rpath,<(module_root_dir)/../../webgpu_path_tracer_b200"],
    "cflags_cc": ["-std=c++17"]
  }]
}
