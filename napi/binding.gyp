{
  "targets": [{
    "target_name": "sail_napi",
    "sources": ["sail_napi.cc"],
    "include_dirs": ["../include"],
    "libraries": ["-L<(module_root_dir)/../sail_b200", "-lsail_b200", "-Wl,-rpath,<(module_root_dir)/../sail_b200"]
  }]
}
