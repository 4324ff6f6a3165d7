{
  "targets": [{
    "target_name": "luxfhe_b200",
    "sources": ["luxfhe_b200_napi.c"],
    "include_dirs": ["../../include"],
    "defines": ["NAPI_VERSION=6"],
    "libraries": ["-L<(module_root_dir)/../../tfhe_b200", "-l:libluxtfhe_b200.so", "-Wl,-rpath,<(module_root_dir)/../../tfhe_b200"]
  }]
}
