#!/bin/bash
echo default; python tools/fill_only.py 2>&1 | tail -2
echo fake; KC_FILL_FAKE=1 python tools/fill_only.py 2>&1 | tail -2
echo fake+rec32; KC_FILL_FAKE=1 KC_REC_MIN_WORDS=8 python tools/fill_only.py 13333334 21 2>&1 | tail -1
echo rec32; KC_REC_MIN_WORDS=8 python tools/fill_only.py 13333334 21 2>&1 | tail -1
