"""vulkan-pathtracer_b200 — B200-native backend for the trace loop of antlilja/vulkan-pathtracer.

The directory name is not a valid Python identifier; load the host binding with
`sys.path.insert(0, "<repo>/vulkan-pathtracer_b200"); import b200pt`.
"""
