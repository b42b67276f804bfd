#pragma once
// placeholder — bit_vector_{stat,sd,rrr} device layouts (K5)
