from .constants import *


def get_aio_config(param_dict):
    a = (param_dict or {}).get(AIO, {}) or {}
    return {
        AIO_BLOCK_SIZE: a.get(AIO_BLOCK_SIZE, AIO_BLOCK_SIZE_DEFAULT),
        AIO_QUEUE_DEPTH: a.get(AIO_QUEUE_DEPTH, AIO_QUEUE_DEPTH_DEFAULT),
        AIO_THREAD_COUNT: a.get(AIO_THREAD_COUNT, AIO_THREAD_COUNT_DEFAULT),
        AIO_SINGLE_SUBMIT: a.get(AIO_SINGLE_SUBMIT, AIO_SINGLE_SUBMIT_DEFAULT),
        AIO_OVERLAP_EVENTS: a.get(AIO_OVERLAP_EVENTS, AIO_OVERLAP_EVENTS_DEFAULT),
    }
