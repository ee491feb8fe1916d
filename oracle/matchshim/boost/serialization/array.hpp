// oracle/matchshim: empty stand-in for <boost/serialization/array.hpp> (test infrastructure); names come from facade.h
