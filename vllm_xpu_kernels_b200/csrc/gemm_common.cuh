// placeholder, filled in with the tcgen05 / TMA helpers
#pragma once
