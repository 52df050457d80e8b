#!/bin/bash
env AVC_LSTM_TRACE=1 PW=1 PK=1 timeout 300 python tools/profile_step.py 2>&1 | grep -A17 "ll2_lstm trace" | tail -38
