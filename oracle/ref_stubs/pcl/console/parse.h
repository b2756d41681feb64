// TEST INFRASTRUCTURE: stand-in (nothing of pcl::console is used by the hot-path files).
