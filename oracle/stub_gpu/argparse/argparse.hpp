/* Empty stand-in: the GPU sources never touch the argument parser. */
#pragma once
namespace argparse { class ArgumentParser {}; }
