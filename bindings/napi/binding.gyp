{
  "targets": [
    {
      "target_name": "g16b200",
      "sources": ["g16_napi.cc"],
      "include_dirs": ["<!@(node -p \"require('node-addon-api').include\")", "../../include"],
      "defines": ["NAPI_DISABLE_CPP_EXCEPTIONS"],
      "libraries": ["-L<(module_root_dir)/../../zk-proof-of-assets_b200", "-lg16b200",
                    "-Wl,-rpath,<(module_root_dir)/../../zk-proof-of-assets_b200"],
      "cflags_cc": ["-std=c++17"]
    }
  ]
}
