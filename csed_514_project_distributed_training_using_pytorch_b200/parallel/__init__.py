from .process_group import (init_process_group, destroy_process_group, get_rank,  # noqa: F401
                            get_world_size, get_comm, is_initialized, send, recv, broadcast,
                            all_reduce, barrier)
from .launch import launch  # noqa: F401
