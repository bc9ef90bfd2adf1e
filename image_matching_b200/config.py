"""Constants of /root/reference/include/config.h (values must match)."""
MATCH_THRESHOLD = 0.44      # config.h:9
COMP_DEPTH = 10             # config.h:14
ALPHA_DEPTH = 2             # config.h:18
MAX_NUM_CORES = 48          # config.h:21 (host threads; unused on the device path)
VECTOR_DIM = 512            # config.h:30
CHUNK_LEN = 128             # config.h:34
