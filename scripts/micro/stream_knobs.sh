for k in 0 1 2 4 6 7; do echo "== DBG $k"; DQNB_STREAM_DBG=$k python scripts/profile_update.py tf32 4096 2>&1 | sed -n 2,17p | cut -c1-32; done
